#!/bin/bash
# Multi-GPU evidence on one box: both bench arms on N GPUs (replicas + the partitioned batch of
# BASELINE configs[4]).  usage: bash tools/r2_multi.sh N [tag]
N=${1:-8}
TAG=${2:-r2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29511 bench.py --gpus $N --steps 50 --warmup 5 --no-configs > gpurun_out/${TAG}_bench_ours_${N}gpu.json 2> gpurun_out/${TAG}_m1.err
tail -2 gpurun_out/${TAG}_m1.err
$TR --master-port 29512 bench.py --impl reference --gpus $N --steps 20 --warmup 3 --no-configs > gpurun_out/${TAG}_bench_reference_${N}gpu.json 2> gpurun_out/${TAG}_m2.err
tail -2 gpurun_out/${TAG}_m2.err
python tools/show_bench.py gpurun_out/${TAG}_bench_ours_${N}gpu.json gpurun_out/${TAG}_bench_reference_${N}gpu.json
echo host cores: $(nproc)
$TR --master-port 29513 tools/batch_scale.py 4096 128:3 64:4 32:4 > gpurun_out/${TAG}_batch_scale_${N}gpu.txt 2> gpurun_out/${TAG}_m3.err
cat gpurun_out/${TAG}_batch_scale_${N}gpu.txt; tail -2 gpurun_out/${TAG}_m3.err
