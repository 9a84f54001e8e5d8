#!/bin/bash
# One profiling gpurun call (1 GPU): GPU tests, bench lines, reference arm, ncu launch lists and full captures.
#   gpurun --timeout 2400 -- 'bash tools/gpu_call_profile.sh r2a'
# Everything lands in gpurun_out/ with the tag; tools/make_profiles.py <tag> turns it into profiles/.
TAG=${1:-r2a}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > $O/smi_$TAG.txt 2>&1
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $O/pytest_$TAG.log 2>&1
echo "pytest rc=$?" >> $O/pytest_$TAG.log
tail -3 $O/pytest_$TAG.log
timeout 600 python bench.py --steps 100 --warmup 5 > $O/bench_$TAG.json 2> $O/bench_$TAG.err
echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 2 > $O/bench_${TAG}_ref.json 2> $O/bench_${TAG}_ref.err
echo "ref rc=$?"
timeout 300 python bench.py --workload c3 --steps 50 --warmup 5 > $O/bench_${TAG}_c3.json 2> $O/bench_${TAG}_c3.err
timeout 300 python bench.py --workload c5 --steps 50 --warmup 5 > $O/bench_${TAG}_c5.json 2> $O/bench_${TAG}_c5.err
timeout 300 python bench.py --workload c4n30 --steps 50 --warmup 5 --no-cpu-baseline > $O/bench_${TAG}_n30.json 2> $O/bench_${TAG}_n30.err
for w in c5x4 c5x8; do     # capacity scenes (16384 / 32768 vehicles), with the CPU-twin comparison
  timeout 600 python bench.py --workload $w --steps 50 --warmup 5 > $O/bench_${TAG}_${w}.json 2> $O/bench_${TAG}_${w}.err
  timeout 300 python tools/coop_phases.py $w > $O/coop_phases_${w}_${TAG}.txt 2>&1
done
timeout 300 python tools/coop_phases.py c5 > $O/coop_phases_c5_${TAG}.txt 2>&1
timeout 300 python tools/coop_phases.py c3 > $O/coop_phases_c3_${TAG}.txt 2>&1
echo "coupled benches done"
# launch lists (cold-cache, serialised: shares only)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_$TAG.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-coupled > $O/ncu_l_$TAG.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_${TAG}_c5.csv \
    python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_l5_$TAG.log 2>&1
# full captures of the two dominant kernels
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 3 -c 1 -f -o $O/prof_$TAG \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-coupled > $O/ncu_f_$TAG.log 2>&1
echo "ncu fused rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_coop -s 3 -c 1 -f -o $O/prof_${TAG}_coop \
    python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > $O/ncu_f5_$TAG.log 2>&1
echo "ncu coop rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_joint -s 3 -c 1 -f -o $O/prof_${TAG}_joint \
    python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/ncu_fj_$TAG.log 2>&1
echo "ncu joint rc=$?"
ls -la $O | tail -30
