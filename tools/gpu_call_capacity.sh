#!/bin/bash
# One-GPU call for the capacity scenes (c5x4 = 16384, c5x8 = 32768 vehicles): bench record with CPU-twin parity, phase timeline, launch list.
O=gpurun_out
mkdir -p $O
for w in c5x4 c5x8; do
  timeout 600 python bench.py --workload $w --steps 20 --warmup 3 > $O/bench_${w}_1gpu.json 2> $O/bench_${w}_1gpu.err
  echo "$w rc=$?"; tail -c 400 $O/bench_${w}_1gpu.err
  timeout 300 python tools/coop_phases.py $w > $O/coop_phases_${w}_1gpu.txt 2>&1
  tail -6 $O/coop_phases_${w}_1gpu.txt
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches_c5x4.csv python bench.py --workload c5x4 --steps 2 --warmup 1 --no-cpu-baseline > /dev/null 2>&1
grep -c k_ $O/launches_c5x4.csv
