#!/usr/bin/env bash
# Runs on the GPU box (via gpurun): tests, smoke, bench, then the ncu passes.
# Usage: tools/gpu_round.sh <tag> [skip-tests]
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
if [ "${2:-}" != "skip-tests" ]; then
  timeout 900 python -m pytest tests -x -q -m gpu > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
  tail -5 $OUT/pytest_gpu_$TAG.log
  timeout 300 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke_$TAG.log
fi
timeout 900 python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cat $OUT/bench_$TAG.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-configs"
timeout 300 $CMD > $OUT/plain_$TAG.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launch_$TAG.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $CMD > $OUT/plain2_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:backtrace_kernel -s 3 -c 1 -f -o $OUT/prof_adj_$TAG $CMD > $OUT/ncu_full_adj_$TAG.log 2>&1
echo "ncu full adj rc=$?"
timeout 300 $CMD > $OUT/plain3_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:^trace_kernel -s 3 -c 1 -f -o $OUT/prof_fwd_$TAG $CMD > $OUT/ncu_full_fwd_$TAG.log 2>&1
echo "ncu full fwd rc=$?"
ls -la $OUT | tail -20
