#!/usr/bin/env python
"""Where a fused batch spends its time (one GPU): host planning, host enqueue, device time per chunk
for several chunk sizes.  usage: batch_probe.py [images=512]"""
import ctypes as C
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np
import torch

import bench
from jpeggpu_b200 import BatchItem
from jpeggpu_b200.batch import FusedBatch, make_items


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    paths = bench.batch_paths(list(range(n)))
    blobs = [open(p, "rb").read() for p in paths]
    offs, at = [], 0
    for b in blobs:
        offs.append(at)
        at += (len(b) + 15) & ~15
    host = torch.empty(at, dtype=torch.uint8, pin_memory=True)
    hv = host.numpy()
    for o, b in zip(offs, blobs):
        hv[o:o + len(b)] = np.frombuffer(b, dtype=np.uint8)
    items = make_items([(host.data_ptr() + o, len(b)) for o, b in zip(offs, blobs)])
    isz = C.sizeof(BatchItem)
    dev = torch.device("cuda:0")
    stream = torch.cuda.current_stream().cuda_stream
    for chunk in (16, 32, 64, 128, 256):
        if chunk > n:
            break
        fb = FusedBatch()
        fb.set_host_arena(host.data_ptr(), host.numel())
        sub = C.cast(C.addressof(items), C.c_void_p)
        t0 = time.perf_counter()
        reps = 20
        for _ in range(reps):
            need, _ = fb.plan(sub, count=chunk)
            fb.lib.jpeggpu_b200_batch_plan  # noqa
        # (the loop above waits for nothing: no decode has been enqueued yet)
        plan_us = (time.perf_counter() - t0) / reps / chunk * 1e6
        arena = torch.empty(need, dtype=torch.uint8, device=dev)
        planes = [[torch.empty((items[j].info.sizes_y[c], items[j].info.sizes_x[c]), dtype=torch.uint8, device=dev)
                   for c in range(items[j].info.num_components)] for j in range(chunk)]
        for j in range(chunk):
            for c, p in enumerate(planes[j]):
                items[j].img.image[c] = p.data_ptr()
                items[j].img.pitch[c] = p.stride(0)
        fb.decode(sub, arena.data_ptr(), arena.numel(), stream, count=chunk)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        enq = 0.0
        a.record()
        for _ in range(reps):
            fb.plan(sub, count=chunk)
            h0 = time.perf_counter()
            fb.decode(sub, arena.data_ptr(), arena.numel(), stream, count=chunk)
            enq += time.perf_counter() - h0
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / reps
        print(f"chunk {chunk:4d}: plan {plan_us:6.2f} us/image  decode-call {enq / reps / chunk * 1e6:6.2f} us/image  "
              f"device (upload + 5 kernels, serial batches) {ms:7.3f} ms/chunk = {ms / chunk * 1e3:6.2f} us/image "
              f"= {chunk / ms:7.1f} k images/s  arena {need / chunk / 1e6:.2f} MB/image")
        fb.close()


if __name__ == "__main__":
    main()
