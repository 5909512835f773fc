#!/usr/bin/env python
"""Where the kernels of one decode sit in time (PDL overlaps them, so per-kernel durations do not
add up): globaltimer stamps written by the diagnostic build (tools/build_timeline.sh,
JPEGGPU_B200_LIB is set to it here).  All times in microseconds from the first destuff CTA.
usage: timeline.py [files...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.environ["JPEGGPU_B200_LIB"] = os.path.join(ROOT, "jpeggpu_b200", "lib", "diag", "lib_timeline.so")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
import torch

import gen_inputs
from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes

# slot -> (label, stored inverted)
SLOTS = {0: ("destuff first CTA past the dependency wait", True), 1: ("destuff last CTA end", False),
         2: ("sync first CTA start (prologue: tables, zero-fill)", True),
         20: ("sync last CTA done with its prologue", False),
         21: ("sync last fill warp done (zero-fill of coefficients + DC)", False),
         3: ("sync first CTA past the wait", True), 4: ("sync last CTA past the wait", False),
         13: ("sync last window staged", False), 5: ("sync last CTA done with its chunks", False),
         14: ("sync verify done (last CTA)", False), 6: ("sync end", False),
         7: ("write first CTA start (prologue)", True), 15: ("write first CTA past the wait", True),
         8: ("write last CTA past the wait", False), 9: ("write end", False),
         16: ("destuff last ticket taken", False), 18: ("destuff first aggregate published", True),
         17: ("destuff last aggregate published", False), 19: ("destuff last look-back done", False),
         10: ("dc first CTA start", True), 11: ("dc first CTA past the wait", True), 12: ("dc end", False)}
ORDER = [0, 16, 18, 17, 19, 2, 20, 1, 3, 4, 13, 7, 21, 5, 14, 6, 15, 8, 10, 9, 11, 12]


def main():
    files = sys.argv[1:] or [gen_inputs.ensure_synth(4032, 3024, 6510), gen_inputs.ensure_img6510()]
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda:0")
    for path in files:
        data = open(path, "rb").read()
        sess = DecodeSession("cuda:0")
        host = pinned_bytes(data)
        sess.parse(host)
        planes = sess.alloc_planes()
        sess.transfer()
        st = sess.dec.scan_stats(0)
        rows = []
        for rep in range(7):
            sess.decode(planes)
            torch.cuda.synchronize()
            sess.scratch[st.off_control:st.off_control + 128].zero_()
            flush.fill_(rep)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            sess.decode(planes)
            b.record()
            torch.cuda.synchronize()
            tl = sess.read_scratch(st.off_control, 128).view(np.uint32).astype(np.int64)
            t0 = (~tl[0]) & 0xffffffff
            row = {k: ((((~tl[k]) & 0xffffffff) if inv else tl[k]) - t0) / 1e3 for k, (_, inv) in SLOTS.items()}
            row["total"] = a.elapsed_time(b) * 1e3
            rows.append(row)
        rows.sort(key=lambda r: r["total"])
        med = rows[len(rows) // 2]
        print(f"== {os.path.basename(path)}: decode {med['total']:.1f} us (median of {len(rows)}, L2 flushed)")
        for k in ORDER:
            print(f"   {med[k]:8.2f}  {SLOTS[k][0]}")
        sess.close()


if __name__ == "__main__":
    main()
