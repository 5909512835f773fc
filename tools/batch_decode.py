#!/usr/bin/env python
"""BASELINE config 5: a batch of synthetic 1920x1080 4:2:0 q90 stb JPEGs partitioned across the
GPUs of one box as independent per-GPU streams (no NCCL on the data path), aggregate images/s.

    python tools/batch_decode.py --images 4096 --distinct 64                       # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29511 tools/batch_decode.py --images 4096 --distinct 64      # 8 GPUs

`--distinct K` distinct files (seeds 0..K-1, written once by oracle/_ref/stbgen into data/_gen)
are cycled to `--images` items; generating 4096 distinct files takes ~10 min of CPU and changes
nothing for the decoder (no state is carried between images except the cached Huffman LUTs, which
every stb file shares anyway).  Prints one JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import time

# More lanes than hardware work queues (8 by default) makes streams share a queue and wait for
# each other: 2 MP batch, 1 GPU, 24 lanes: 35.1 k images/s with 8 queues, 40.9 k with 32
# (profiles/r1_v8_batch_connections.txt).  Must be set before the CUDA context exists.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--images", type=int, default=4096)
    ap.add_argument("--distinct", type=int, default=32)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--lanes", type=int, default=4)
    ap.add_argument("--repeats", type=int, default=3)
    ap.add_argument("--verify", action="store_true", help="check a sample against the oracle")
    args = ap.parse_args()

    import torch
    import gen_inputs
    from jpeggpu_b200 import batch as jb
    from jpeggpu_b200.torch_io import pinned_bytes

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group(backend="nccl", device_id=torch.device(f"cuda:{local}"))

    def barrier():
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    paths = [gen_inputs.ensure_synth(args.width, args.height, seed, 90) for seed in range(args.distinct)]
    if any(p is None for p in paths):
        raise SystemExit("synthetic inputs missing and oracle/_ref/stbgen not built")
    blobs = [open(p, "rb").read() for p in paths]
    mine = jb.partition(args.images, rank, world)
    pinned = [pinned_bytes(b) for b in blobs]
    files = [pinned[i % args.distinct] for i in mine]

    bd = jb.BatchDecoder(f"cuda:{local}", lanes=args.lanes)
    bd.prepare(files, plane_sets=4 * args.lanes)
    bd.run()
    bd.synchronize()  # warm-up (kernel attributes, LUT cache, pinned staging)

    if args.verify and rank == 0:
        import numpy as np
        import oracle_bindings as ob
        k = min(len(files), len(bd.planes)) - 1
        # the last writer of plane set k is the last item i with i % sets == k
        sets = len(bd.planes)
        last = max(i for i in range(len(files)) if i % sets == k)
        ref = ob.decode(blobs[mine[last] % args.distinct])
        for c, p in enumerate(bd.planes[k]):
            assert np.array_equal(p.cpu().numpy(), ref["pixels"][c]), "batch output differs from oracle"

    best = None
    for _ in range(args.repeats):
        barrier()
        t0 = time.perf_counter()
        bd.run()
        bd.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            import torch.distributed as dist
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        best = dt if best is None else min(best, dt)
    counts = jb.gather_counts({"rank": rank, "items": len(mine)}, world)
    if rank == 0:
        total = sum(c["items"] for c in counts)
        mp = args.width * args.height / 1e6
        print(json.dumps({
            "metric": "batch decode throughput", "value": round(total / best, 1), "unit": "images/s",
            "mp_per_s": round(total * mp / best, 1), "n_gpus": world, "images": total,
            "seconds": round(best, 5), "scaling": "strong",
            "config": {"workload": f"{total} x {args.width}x{args.height} 4:2:0 q90 stb "
                                   f"({args.distinct} distinct files cycled), {args.lanes} streams/GPU, "
                                   "host input pinned, parse+H2D+decode per image, wall clock "
                                   "max over ranks, best of %d" % args.repeats,
                       "parallelism": f"batch partitioned round-robin over {world} GPU(s), no collective"},
            "per_rank_items": [c["items"] for c in counts]}))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
