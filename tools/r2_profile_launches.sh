#!/bin/bash
# ncu launch list of the bench command (B200_PROFILING.md: per-launch times are cold-cache and
# serialised; the kernels' SHARES of the step are what must agree with bench.py).  One B200.
set -u
TAG=${1:-r2}
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-configs --batch-images 0"
$CMD > gpurun_out/${TAG}_ncu_bench_plain.log 2>&1 || { tail -5 gpurun_out/${TAG}_ncu_bench_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
  --log-file gpurun_out/${TAG}_ncu_launches.csv $CMD > gpurun_out/${TAG}_ncu_bench.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_bench.log; wc -l gpurun_out/${TAG}_ncu_launches.csv
