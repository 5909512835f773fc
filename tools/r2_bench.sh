#!/bin/bash
# Round-2 bench run (one B200): GPU tests, both bench arms with all legs.  Outputs gpurun_out/<tag>_*.
set -u
TAG=${1:-r2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/${TAG}_pytest_gpu.txt 2>&1; tail -3 gpurun_out/${TAG}_pytest_gpu.txt
( time python bench.py --steps 50 --warmup 5 > gpurun_out/${TAG}_bench_ours.json 2> gpurun_out/${TAG}_bench_ours.err ) 2>&1 | grep real
tail -3 gpurun_out/${TAG}_bench_ours.err
( time python bench.py --impl reference --steps 50 --warmup 5 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err ) 2>&1 | grep real
tail -3 gpurun_out/${TAG}_bench_reference.err
python tools/show_bench.py gpurun_out/${TAG}_bench_ours.json gpurun_out/${TAG}_bench_reference.json
echo done
