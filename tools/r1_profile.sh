#!/bin/bash
# Round-1 evidence run (one B200): tests, both bench arms, ncu launch list, ncu full captures.
# Usage on the GPU box: bash tools/r1_profile.sh   (outputs under gpurun_out/)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest_gpu.log 2>&1; tail -2 gpurun_out/r1_pytest_gpu.log
python bench.py --impl reference --steps 50 --warmup 5 > gpurun_out/r1_bench_reference.json 2> gpurun_out/r1_bench_reference.err
python bench.py --steps 100 --warmup 5 > gpurun_out/r1_bench_ours.json 2> gpurun_out/r1_bench_ours.err
rc=$?
python tools/kernel_times.py > gpurun_out/r1_kernel_times.txt 2>&1
python tools/ctrl_dump.py > gpurun_out/r1_ctrl.txt 2>&1
python tools/e2e_breakdown.py synth > gpurun_out/r1_e2e_breakdown.txt 2>&1
python tools/e2e_breakdown.py img >> gpurun_out/r1_e2e_breakdown.txt 2>&1
if [ $rc -eq 0 ]; then
  # launch list of the bench command (B200_PROFILING.md): per-launch times are cold-cache and serialised
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/r1_ncu_launches.csv python bench.py --steps 3 --warmup 3 > gpurun_out/r1_ncu_bench.log 2>&1
  python tools/profile_decode.py > gpurun_out/r1_pd.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on --launch-skip 5 --launch-count 5 \
    -o gpurun_out/r1_full -f python tools/profile_decode.py > gpurun_out/r1_ncu_full.log 2>&1
fi
echo done rc=$rc
