#!/bin/bash
# One-GPU call after a change to the coupled path: full GPU test suite, then the coupled records (BASELINE configs 3 and 5
# and the two capacity scenes) with the CPU-twin comparison, phase timelines.   gpurun -- 'bash tools/gpu_call_coupled_ab.sh <tag>'
TAG=${1:-ab}
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_${TAG}.log 2>&1
echo "pytest rc=$?"; tail -3 $O/pytest_${TAG}.log
for w in c5 c3 c5x4 c5x8; do
  timeout 600 python bench.py --workload $w --steps 50 --warmup 5 > $O/bench_${w}_${TAG}.json 2> $O/bench_${w}_${TAG}.err
  echo "$w rc=$?"
  python - <<PY
import json
r = json.loads(open("$O/bench_${w}_${TAG}.json").read().strip().split("\n")[-1])
print("$w", r["ms_per_step"], r["e2e"]["ms_per_step"], r.get("parity_vs_cpu_port"), r["gpu_launches"])
PY
  timeout 300 python tools/coop_phases.py $w > $O/coop_phases_${w}_${TAG}.txt 2>&1
  head -6 $O/coop_phases_${w}_${TAG}.txt
done
