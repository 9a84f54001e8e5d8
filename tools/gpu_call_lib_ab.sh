#!/bin/bash
# A/B of an alternative build of the library on the coupled records:  gpurun -- 'bash tools/gpu_call_lib_ab.sh <lib.so> <tag>'
LIB=$1
TAG=${2:-ab}
O=gpurun_out
mkdir -p $O
for w in c5 c3 c5x4 c5x8; do
  for v in base alt; do
    if [ $v = alt ]; then export PIADMM_LIB=$LIB; else unset PIADMM_LIB; fi
    timeout 300 python bench.py --workload $w --steps 50 --warmup 5 --no-cpu-baseline > $O/bench_${w}_${TAG}_$v.json 2> $O/bench_${w}_${TAG}_$v.err
    python - <<PY
import json
r = json.loads(open("$O/bench_${w}_${TAG}_$v.json").read().strip().split("\n")[-1])
print("$w", "$v", round(r["ms_per_step"], 4), round(r["e2e"]["ms_per_step"], 4), r["result_sha256"][:12])
PY
  done
done
export PIADMM_LIB=$LIB
timeout 300 python tools/coop_phases.py c5 2>&1 | head -6
