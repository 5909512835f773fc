#!/bin/bash
# Multi-GPU evidence on one box: bench.py replicas (ours + reference arm) and the partitioned
# batch (BASELINE config 5) on N GPUs.  usage: bash tools/r1_multi.sh N
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29511 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/r1_bench_ours_${N}gpu.json 2> gpurun_out/m1.err
$TR --master-port 29512 bench.py --impl reference --gpus $N --steps 20 --warmup 3 > gpurun_out/r1_bench_reference_${N}gpu.json 2> gpurun_out/m2.err
for l in 8 16; do
  $TR --master-port $((29520 + l)) tools/batch_decode.py --images 16384 --distinct 32 --lanes $l 2> gpurun_out/m3.err | tail -1 > gpurun_out/r1_batch_${N}gpu_lanes$l.json
done
python - <<P
import json
for f in ["r1_bench_ours_${N}gpu.json", "r1_batch_${N}gpu_lanes8.json", "r1_batch_${N}gpu_lanes16.json"]:
    try:
        d = json.load(open("gpurun_out/" + f)); print(f, d["n_gpus"], d["value"], d.get("e2e", {}).get("value"))
    except Exception as e:
        print(f, "ERR", e)
P
echo host cores: $(nproc)
