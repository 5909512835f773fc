#!/bin/bash
# Evidence run for the RGB output kernels (one B200): GPU tests, bench line, route timings, ncu.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest_gpu.log 2>&1; tail -2 gpurun_out/r1_pytest_gpu.log
python bench.py --steps 100 --warmup 5 > gpurun_out/r1_bench_ours.json 2> gpurun_out/r1_bench_ours.err; echo bench rc=$?
python tools/rgbi_times.py > gpurun_out/r1_rgbi_times.txt 2>&1
python tools/kernel_times.py > gpurun_out/r1_kernel_times.txt 2>&1
python tools/profile_rgbi.py > gpurun_out/r1_prgbi.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'rgbi|idct_kernel' --launch-skip 4 --launch-count 3 \
  -o gpurun_out/r1_rgbi_full -f python tools/profile_rgbi.py > gpurun_out/r1_ncu_rgbi.log 2>&1
tail -3 gpurun_out/r1_ncu_rgbi.log
echo done
