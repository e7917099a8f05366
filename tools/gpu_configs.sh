#!/bin/bash
# BASELINE configs C1-C3, the Sponza camera sweep and the shift-mapped mode through bench.py (one JSON line each).
# Usage: tools/gpu_configs.sh TAG   ->  gpurun_out/configs_TAG.jsonl
TAG=${1:-x}
O=gpurun_out/configs_$TAG.jsonl
: > $O
run() { echo "== $*" >&2; python bench.py "$@" 2>/dev/null >> $O || echo "{\"failed\": \"$*\"}" >> $O; }
run --scene cbox_gdpt --steps 20 --warmup 5 --cpu-seconds 5                      # C1: 512x512, 4 spp
run --scene veach_mi_gradpath --steps 10 --warmup 3 --cpu-seconds 8              # C2: 768x512, 64 spp
run --scene disney_gradpath --steps 5 --warmup 3 --cpu-seconds 10               # C3: 683x512, 256 spp
run --scene sponza_1080p --spp 64 --steps 3 --warmup 3 --no-cpu-baseline --f64-steps 0                # camera 0 at the sweep's spp
run --scene sponza_1080p --spp 64 --camera 1 --steps 3 --warmup 3 --no-cpu-baseline --f64-steps 0
run --scene sponza_1080p --spp 64 --camera 2 --steps 3 --warmup 3 --no-cpu-baseline --f64-steps 0
run --scene sponza_1080p --spp 64 --offsets shift --steps 3 --warmup 3 --no-cpu-baseline --f64-steps 0   # the opt-in shift-mapped mode
python - <<PY
import json
for l in open("$O"):
    d = json.loads(l)
    if "failed" in d: print(d); continue
    print("%-90s %9.1f Mpaths/s  e2e %9.1f  %s" % (d["config"]["workload"], d["value"], d["e2e"]["value"], ("cpu %.2f" % d["cpu_baseline"]["value"]) if d.get("cpu_baseline", {}).get("value") else ""))
PY
