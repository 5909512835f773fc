#!/bin/bash
# One `ncu --set full` capture of the five kernels of a decode (one B200).  The command runs once
# without ncu first.  usage: bash tools/r2_profile_full.sh [tag] [file]
set -u
TAG=${1:-r2}
FILE=${2:-data/_gen/synth_4032x3024_s6510_q90.jpg}
mkdir -p gpurun_out
python tools/profile_decode.py $FILE 3 > gpurun_out/${TAG}_pd.log 2>&1 || { cat gpurun_out/${TAG}_pd.log; exit 1; }
# 3 decodes x 5 kernels: skip the first two decodes, capture the third
timeout 900 ncu --set full --clock-control none --import-source on --launch-skip 10 --launch-count 5 \
  -o gpurun_out/${TAG}_full -f python tools/profile_decode.py $FILE 3 > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -3 gpurun_out/${TAG}_ncu_full.log
python tools/ncu_summary.py gpurun_out/${TAG}_full.ncu-rep gpurun_out/${TAG}_ncu_full_summary.json && cat gpurun_out/${TAG}_ncu_full_summary.json | head -c 3000
