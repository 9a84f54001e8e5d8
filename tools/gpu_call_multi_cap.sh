#!/bin/bash
# Multi-GPU call for the coupled path:  gpurun --gpus N -- 'bash tools/gpu_call_multi_cap.sh N <tag> [full]'
# The multi-rank pytest cases (or the whole GPU suite with `full`), the vehicle-sharded check against the CPU twin (peer and
# NCCL exchange), and the coupled records of BASELINE config 5 and the capacity scenes at N ranks, with phase timelines.
N=${1:-2}
TAG=${2:-r2cap}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$3" = "full" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_${TAG}_${N}gpu.log 2>&1
else
  timeout 600 python -m pytest tests/test_gpu_coupled.py -m gpu -x -q -k "two_ranks or sharded" > $O/pytest_${TAG}_${N}gpu.log 2>&1
fi
echo "pytest rc=$?"; tail -3 $O/pytest_${TAG}_${N}gpu.log
timeout 600 $TR --master-port 29502 tests/mgpu_c5.py > $O/mgpu_c5_${TAG}_${N}gpu.jsonl 2> $O/mgpu_c5_${TAG}_${N}gpu.err
echo "mgpu_c5 rc=$?"
port=29510
for w in c5 c5x4 c5x8; do
  port=$((port+1))
  timeout 600 $TR --master-port $port bench.py --workload $w --gpus $N --steps 50 --warmup 5 > $O/bench_${w}_${TAG}_${N}gpu.json 2> $O/bench_${w}_${TAG}_${N}gpu.err
  echo "$w rc=$?"
  python - <<PY
import json
try:
    r = json.loads(open("$O/bench_${w}_${TAG}_${N}gpu.json").read().strip().split("\n")[-1])
    print("$w", r["ms_per_step"], r["e2e"]["ms_per_step"], r.get("exchange"), r.get("result_sha256", "")[:12], r["work_per_step"]["outer_iters"])
except Exception as e:
    print("$w ERR", e)
PY
  port=$((port+1))
  timeout 300 $TR --master-port $port tools/coop_phases.py $w > $O/coop_phases_${w}_${TAG}_${N}gpu.txt 2>&1
  grep -A5 "^rank 0" $O/coop_phases_${w}_${TAG}_${N}gpu.txt | head -6
done
