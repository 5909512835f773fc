#!/usr/bin/env python
"""Host-side breakdown of the e2e protocol on the bench workload: microseconds spent in each C-ABI
call and waiting for the device (wall clock, median of N)."""
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import gen_inputs  # noqa: E402
from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "synth"
    if name == "synth":
        path = gen_inputs.ensure_synth(4032, 3024, 6510, 90)
    elif name == "img":
        path = gen_inputs.ensure_img6510()
    else:
        path = name
    data = open(path, "rb").read()
    sess = DecodeSession("cuda:0")
    if os.environ.get("JB2_EAGER_WALK") == "1":  # A/B of jpeggpu_b200_set_deferred_walk
        sess.dec.set_deferred_walk(False)
    elif os.environ.get("JB2_WALK_THREADS") == "0":  # deferred walk on the calling thread alone
        sess.dec.set_deferred_walk(2)
    host = pinned_bytes(data)
    sess.parse(host)
    planes = sess.alloc_planes()
    sample = torch.empty(4096, dtype=torch.uint8, pin_memory=True)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    rows = {k: [] for k in ("parse", "transfer", "decode", "d2h_enqueue", "sync", "total")}
    for it in range(60):
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sess.parse(host)
        t1 = time.perf_counter()
        sess.transfer()
        t2 = time.perf_counter()
        sess.decode(planes)
        t3 = time.perf_counter()
        sample.copy_(planes[0].view(-1)[:4096], non_blocking=True)
        t4 = time.perf_counter()
        torch.cuda.synchronize()
        t5 = time.perf_counter()
        if it >= 10:
            for k, v in zip(rows, (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t5 - t0)):
                rows[k].append(v * 1e6)
    print(os.path.basename(path), " ".join(f"{k}={statistics.median(v):.1f}us" for k, v in rows.items()))


if __name__ == "__main__":
    main()
