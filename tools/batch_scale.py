#!/usr/bin/env python
"""The batch leg of bench.py alone (BASELINE configs[4]: 4096 distinct 1920x1080 files, file i on
rank i mod N), for several batch sizes / depths.  Run directly (1 GPU) or under torchrun.
usage: batch_scale.py [images=4096] [chunk:depth ...]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    rank, world, local = bench.dist_setup()
    device = torch.device(f"cuda:{local}")
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    combos = [tuple(int(x) for x in a.split(":")) for a in sys.argv[2:]] or [(128, 3), (64, 3), (32, 4)]
    for chunk, depth in combos:
        r = bench.ours_batch(rank, world, device, n, chunk, depth, repeats=3)
        if "unavailable" in r:
            if rank == 0:
                print(json.dumps(r))
            return
        secs = min(bench.max_over_ranks(t, world) for t in r["seconds_all_passes"])
        host = bench.max_over_ranks(r["host_enqueue_seconds"], world)
        best = (secs, host, r["matches_per_image_path"])
        if rank == 0:
            print(json.dumps({"n_gpus": world, "images": n, "chunk": chunk, "depth": depth,
                              "seconds_best_of_3": round(best[0], 5), "images_per_s": round(n / best[0], 1),
                              "host_enqueue_seconds_max_rank": round(best[1], 5), "matches": best[2],
                              "nproc": os.cpu_count()}), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
