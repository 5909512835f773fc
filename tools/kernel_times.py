#!/usr/bin/env python
"""Per-kernel device times of one decode (CUDA events via jpeggpu_b200_set_profiling) for the
library selected by JPEGGPU_B200_LIB.  usage: kernel_times.py [file ...]"""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import torch

import gen_inputs
from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes


def main():
    files = sys.argv[1:] or [gen_inputs.ensure_synth(4032, 3024, 6510), gen_inputs.ensure_img6510()]
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
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
        reps, tot = 10, 0.0
        for _ in range(reps):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            sess.decode(planes)
            b.record()
            torch.cuda.synchronize()
            tot += a.elapsed_time(b)
        sess.dec.set_profiling(True)
        acc = {}
        for _ in range(reps):
            flush.fill_(1)
            sess.decode(planes)
            torch.cuda.synchronize()
            for k, (ms, n) in sess.dec.kernel_times().items():
                acc[k] = acc.get(k, 0.0) + ms
        sess.dec.set_profiling(False)
        h = hashlib.sha256()
        for p in planes:
            h.update(p.cpu().numpy().tobytes())
        st = sess.dec.scan_stats(0)
        print(f"{os.path.basename(os.environ.get('JPEGGPU_B200_LIB', 'default'))} {os.path.basename(path)} "
              f"N={st.num_subsequences} total_us={tot / reps * 1e3:.1f} sha={h.hexdigest()[:10]} "
              + " ".join(f"{k.replace('huff_', '').replace('_kernel', '')}={v / reps * 1e3:.1f}" for k, v in acc.items()))
        sess.close()


if __name__ == "__main__":
    main()
