#!/bin/bash
# Round-2 evidence pass: the Poisson timing probe, a short plain bench run, then (only after it exited 0) the ncu launch list of
# the same command and --set full captures of the hot kernels.  Usage: tools/gpu_round2.sh TAG
TAG=${1:-r2}
O=gpurun_out
python tools/gpu_poisson_bench.py > $O/poisson_$TAG.log 2>&1; cat $O/poisson_$TAG.log
B="python bench.py --steps 1 --warmup 1 --spp 8 --no-cpu-baseline --f64-steps 0"
$B > $O/plain_$TAG.log 2>&1 || { echo "plain bench failed"; tail -5 $O/plain_$TAG.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/launches_$TAG.csv $B > $O/ncu_${TAG}_0.log 2>&1
F="--set full --clock-control none --import-source on -c 1 -f"
ncu $F -k regex:k_trace -s 0 -o $O/prof_${TAG}_trace_primary $B > $O/ncu_${TAG}_1.log 2>&1
ncu $F -k regex:k_trace -s 2 -o $O/prof_${TAG}_trace_bounce $B > $O/ncu_${TAG}_2.log 2>&1
ncu $F -k regex:k_bounce -s 1 -o $O/prof_${TAG}_bounce $B > $O/ncu_${TAG}_3.log 2>&1
ncu $F -k regex:k_dct1_fft -s 3 -o $O/prof_${TAG}_dct_rows $B > $O/ncu_${TAG}_4.log 2>&1
ncu $F -k regex:k_dct1_fft -s 4 -o $O/prof_${TAG}_dct_cols $B > $O/ncu_${TAG}_5.log 2>&1
ls -la $O | grep $TAG
