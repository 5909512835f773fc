#!/usr/bin/env python
"""Decode one image a few times (nothing else) — the command ncu wraps (B200_PROFILING.md)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch

import gen_inputs
from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes


def main():
    path = sys.argv[1] if len(sys.argv) > 1 else gen_inputs.ensure_synth(4032, 3024, 6510)
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    data = open(path, "rb").read()
    sess = DecodeSession("cuda:0")
    host = pinned_bytes(data)
    sess.parse(host)
    planes = sess.alloc_planes()
    sess.transfer()
    for _ in range(reps):
        sess.decode(planes)
    torch.cuda.synchronize()
    sess.close()
    print("decoded", os.path.basename(path), reps, "times")


if __name__ == "__main__":
    main()
