#!/bin/bash
# Round-2 quick check (one B200): GPU tests, per-kernel times and control blocks for the four
# workloads, tuning variants, a short bench line.  Outputs under gpurun_out/<tag>_*.
set -u
TAG=${1:-r2}
mkdir -p gpurun_out
G=data/_gen
FILES="$G/synth_4032x3024_s6510_q90.jpg $G/IMG_6510.JPG $G/synth_8064x6048_s48_q90.jpg $G/synth_1920x1080_s1_q90.jpg"
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.txt 2>&1; tail -3 gpurun_out/${TAG}_pytest_gpu.txt
python tools/kernel_times.py $FILES > gpurun_out/${TAG}_kernel_times.txt 2>&1
for v in jpeggpu_b200/lib/variants/*.so; do
  [ -f "$v" ] && JPEGGPU_B200_LIB=$v python tools/kernel_times.py $G/synth_4032x3024_s6510_q90.jpg $G/IMG_6510.JPG >> gpurun_out/${TAG}_kernel_times.txt 2>&1
done
cat gpurun_out/${TAG}_kernel_times.txt
python tools/ctrl_dump.py $FILES > gpurun_out/${TAG}_ctrl.txt 2>&1; cat gpurun_out/${TAG}_ctrl.txt
python bench.py --steps 50 --warmup 5 > gpurun_out/${TAG}_bench_ours.json 2> gpurun_out/${TAG}_bench_ours.err; cat gpurun_out/${TAG}_bench_ours.json
echo done
