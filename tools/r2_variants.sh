#!/bin/bash
# kernel times of every library under jpeggpu_b200/lib/variants (plus the default) on the four workloads
G=data/_gen
FILES="$G/synth_4032x3024_s6510_q90.jpg $G/IMG_6510.JPG $G/synth_8064x6048_s48_q90.jpg $G/synth_1920x1080_s1_q90.jpg"
mkdir -p gpurun_out
python tools/kernel_times.py $FILES > gpurun_out/${1:-var}_kernel_times.txt 2>&1
for v in jpeggpu_b200/lib/variants/*.so; do
  [ -f "$v" ] && JPEGGPU_B200_LIB=$v python tools/kernel_times.py $FILES >> gpurun_out/${1:-var}_kernel_times.txt 2>&1
done
cat gpurun_out/${1:-var}_kernel_times.txt
