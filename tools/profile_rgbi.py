#!/usr/bin/env python
"""Runs the two RGB routes a few times (nothing else) - the command ncu wraps for the RGB kernels:
decode_rgbi (idct_rgbi_kernel) and decode + convert_rgbi (ycc_to_rgbi_kernel)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools")]
import torch

import gen_inputs
import jpeggpu_b200 as jb
from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes


def main():
    path = sys.argv[1] if len(sys.argv) > 1 else gen_inputs.ensure_synth(4032, 3024, 6510)
    data = open(path, "rb").read()
    sess = DecodeSession("cuda:0")
    info = sess.parse(pinned_bytes(data))
    planes = sess.alloc_planes()
    rgb = torch.empty((info.sizes_y[0], info.sizes_x[0], 3), dtype=torch.uint8, device="cuda")
    sess.transfer()
    s = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        sess.decode_rgbi(rgb)
        sess.decode(planes)
        jb.convert_rgbi([p.data_ptr() for p in planes], [p.stride(0) for p in planes], info,
                        rgb.data_ptr(), rgb.stride(0), s)
    torch.cuda.synchronize()
    sess.close()
    print("rgb routes done on", os.path.basename(path))


if __name__ == "__main__":
    main()
