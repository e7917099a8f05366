#!/bin/bash
# One GPU visit: parity tests, the Sponza probe, a short plain bench run, then (only after that exited 0)
# the ncu launch list and --set full captures of the hot kernels.  Usage: tools/gpu_round.sh TAG [noprof]
TAG=${1:-x}
O=gpurun_out
python -m pytest tests -m gpu -q > $O/gpu_tests_$TAG.log 2>&1; tail -3 $O/gpu_tests_$TAG.log
python tools/gpu_probe_sponza.py > $O/probe_$TAG.log 2>&1; tail -4 $O/probe_$TAG.log
[ "$2" = noprof ] && exit 0
B="python bench.py --steps 1 --warmup 1 --spp 2 --no-cpu-baseline"
$B > $O/plain_$TAG.log 2>&1 || { echo "plain bench failed"; tail -5 $O/plain_$TAG.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_$TAG.csv $B > $O/ncu_${TAG}_0.log 2>&1
F="--set full --clock-control none --import-source on -c 1 -f"
ncu $F -k regex:k_trace -s 0 -o $O/prof_${TAG}_trace_primary $B > $O/ncu_${TAG}_1.log 2>&1
ncu $F -k regex:k_trace -s 2 -o $O/prof_${TAG}_trace_bounce $B > $O/ncu_${TAG}_2.log 2>&1
ncu $F -k regex:k_bounce -s 1 -o $O/prof_${TAG}_bounce $B > $O/ncu_${TAG}_3.log 2>&1
ls -la $O | grep $TAG
