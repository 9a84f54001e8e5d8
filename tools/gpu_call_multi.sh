#!/bin/bash
# One multi-GPU gpurun call:  gpurun --gpus N --timeout 1500 -- 'bash tools/gpu_call_multi.sh N <tag>'
# Default bench line at N ranks (config 4 weak scaling + the coupled records with the exchange in the data path), the
# reference arm, the vehicle-sharded check against the CPU twin (peer and NCCL), phase timelines, the multi-rank pytest cases.
N=${1:-2}
TAG=${2:-r2m}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > $O/topo_${TAG}_${N}gpu.txt 2>&1
timeout 900 $TR --master-port 29501 bench.py --gpus $N --steps 100 --warmup 5 > $O/bench_${TAG}_${N}gpu.json 2> $O/bench_${TAG}_${N}gpu.err
echo "bench N=$N rc=$?"
timeout 600 $TR --master-port 29502 tests/mgpu_c5.py > $O/mgpu_c5_${TAG}_${N}gpu.jsonl 2> $O/mgpu_c5_${TAG}_${N}gpu.err
echo "mgpu_c5 rc=$?"
timeout 300 $TR --master-port 29503 tools/coop_phases.py c5 > $O/coop_phases_c5_${TAG}_${N}gpu.txt 2>&1
timeout 300 $TR --master-port 29504 tools/coop_phases.py c3 > $O/coop_phases_c3_${TAG}_${N}gpu.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_coupled.py -m gpu -x -q -k "two_ranks" > $O/pytest_${TAG}_${N}gpu.log 2>&1
tail -3 $O/pytest_${TAG}_${N}gpu.log
timeout 600 $TR --master-port 29505 bench.py --impl reference --gpus $N --steps 20 --warmup 2 > $O/bench_${TAG}_${N}gpu_ref.json 2> /dev/null
tail -c 600 $O/mgpu_c5_${TAG}_${N}gpu.jsonl
