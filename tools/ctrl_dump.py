#!/usr/bin/env python
"""Decode an image and print the Huffman stage's control block (diagnostic counters)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
import torch

import gen_inputs
from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes

NAMES = ["relinked", "slow_repairs", "link_rounds", "link_verify_ns", "need_verify", "write_stage_ns",
         "write_decode_ns", "write_flush_ns", "sync_stage_ns", "sync_iter_ns", "reserved1",
         "reserved2", "dc_ticket", "write_ticket", "sync_done"]


def main():
    files = sys.argv[1:] or [gen_inputs.ensure_synth(4032, 3024, 6510), gen_inputs.ensure_img6510()]
    for path in files:
        data = open(path, "rb").read()
        sess = DecodeSession("cuda:0")
        host = pinned_bytes(data)
        sess.parse(host)
        planes = sess.alloc_planes()
        sess.transfer()
        for _ in range(3):
            sess.decode(planes)
        torch.cuda.synchronize()
        st = sess.dec.scan_stats(0)
        ctrl = sess.read_scratch(st.off_control, 64).view(np.uint32)
        print(os.path.basename(path), {n: int(ctrl[i]) for i, n in enumerate(NAMES)})
        sess.close()


if __name__ == "__main__":
    main()
