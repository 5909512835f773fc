#!/usr/bin/env python
"""One fused batch of 2 MP files decoded a few times (nothing else) - the command ncu wraps for the
batched kernels.  usage: profile_batch.py [images=64] [repeats=2]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch

import bench
from jpeggpu_b200.batch import FusedBatch, make_items


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    blobs = [open(p, "rb").read() for p in bench.batch_paths(list(range(n)))]
    offs, at = [], 0
    for b in blobs:
        offs.append(at)
        at += (len(b) + 15) & ~15
    host = torch.empty(at, dtype=torch.uint8, pin_memory=True)
    for o, b in zip(offs, blobs):
        host.numpy()[o:o + len(b)] = np.frombuffer(b, dtype=np.uint8)
    items = make_items([(host.data_ptr() + o, len(b)) for o, b in zip(offs, blobs)])
    fb = FusedBatch()
    fb.set_host_arena(host.data_ptr(), host.numel())
    need, _ = fb.plan(items)
    arena = torch.empty(need, dtype=torch.uint8, device="cuda:0")
    planes = []
    for it in items:
        ps = [torch.empty((it.info.sizes_y[c], it.info.sizes_x[c]), dtype=torch.uint8, device="cuda:0")
              for c in range(it.info.num_components)]
        for c, p in enumerate(ps):
            it.img.image[c] = p.data_ptr()
            it.img.pitch[c] = p.stride(0)
        planes.append(ps)
    stream = torch.cuda.current_stream().cuda_stream
    for r in range(reps):
        if r:
            fb.plan(items)
        fb.decode(items, arena.data_ptr(), arena.numel(), stream)
        torch.cuda.synchronize()
    fb.close()
    print("decoded a batch of", n, "files", reps, "times")


if __name__ == "__main__":
    main()
