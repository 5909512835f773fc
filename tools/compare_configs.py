#!/usr/bin/env python
"""Ours vs the reference CUDA decoder (oracle/_ref/libjpeggpu_ref.so) on the BASELINE configs that
fit one GPU: IMG_6510 (config 0), synthetic 12 MP (1), the 6 MP sampling sweep (2) and synthetic
48 MP without restart markers (3).  Same protocol for both: e2e = parse_header + get_buffer_size
+ transfer + decode + stream sync from pinned host memory (wall clock, median); device = CUDA
events around decode only.  L2 flushed before every iteration.  Prints a markdown table and one
JSON object; run on a B200:  python tools/compare_configs.py > gpurun_out/configs.md"""
import json
import os
import statistics
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import inputs  # noqa: E402
import ref_bindings as rb  # noqa: E402
from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes  # noqa: E402

ITERS = 30


def cases():
    out = [("0: IMG_6510.JPG 4032x3024 4:2:0 DRI", inputs.img6510()),
           ("1: stb 4032x3024 4:2:0 q90", inputs.synth(4032, 3024, 6510, 90)),
           ("2: stb 3008x2000 4:2:0 q90", inputs.synth(3008, 2000, 1, 90)),
           ("2: libjpeg 3008x2000 4:4:4", inputs.pillow_jpeg(w=3008, h=2000, seed=11, mode="YCbCr", subsampling=0)),
           ("2: libjpeg 3008x2000 4:2:2", inputs.pillow_jpeg(w=3008, h=2000, seed=12, mode="YCbCr", subsampling=1)),
           ("2: libjpeg 3008x2000 gray", inputs.pillow_jpeg(w=3008, h=2000, seed=14, mode="L")),
           ("3: stb 8064x6048 4:2:0 q90 no DRI", inputs.synth(8064, 6048, 48, 90)),
           ("5: stb 1920x1080 4:2:0 q90", inputs.synth(1920, 1080, 1, 90))]
    return [(n, d) for n, d in out if d is not None]


def time_ours(data, flush):
    sess = DecodeSession("cuda:0")
    host = pinned_bytes(data)
    info = sess.parse(host)
    planes = sess.alloc_planes()
    e2e, dev = [], []
    for it in range(ITERS + 3):
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sess.parse(host)
        sess.transfer()
        sess.decode(planes)
        torch.cuda.synchronize()
        if it >= 3:
            e2e.append(time.perf_counter() - t0)
    for it in range(ITERS + 3):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        sess.decode(planes)
        b.record()
        torch.cuda.synchronize()
        if it >= 3:
            dev.append(a.elapsed_time(b))
    mp = info.sizes_x[0] * info.sizes_y[0] / 1e6
    sess.close()
    return mp, statistics.median(e2e) * 1e3, statistics.median(dev)


def time_ref(data, flush):
    host = pinned_bytes(data)
    rd = rb.RefDecoder()
    st, info = rd.parse(host.data_ptr(), len(data))
    if st != 0:
        return None, None
    planes = [torch.zeros((info.sizes_y[c], info.sizes_x[c]), dtype=torch.uint8, device="cuda")
              for c in range(info.num_components)]
    need = rd.buffer_size()
    tmp = torch.empty(need, dtype=torch.uint8, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    ptrs = [p.data_ptr() for p in planes]
    pitches = [p.stride(0) for p in planes]
    e2e, dev = [], []
    for it in range(ITERS + 3):
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rd.parse(host.data_ptr(), len(data))
        rd.buffer_size()
        if rd.transfer(tmp.data_ptr(), need, s) != 0 or rd.decode(ptrs, pitches, tmp.data_ptr(), need, s) != 0:
            rd.close()
            return None, None
        torch.cuda.synchronize()
        if it >= 3:
            e2e.append(time.perf_counter() - t0)
    for it in range(ITERS + 3):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rd.decode(ptrs, pitches, tmp.data_ptr(), need, s)
        b.record()
        torch.cuda.synchronize()
        if it >= 3:
            dev.append(a.elapsed_time(b))
    rd.close()
    return statistics.median(e2e) * 1e3, statistics.median(dev)


def main():
    assert rb.available(), "oracle/_ref/libjpeggpu_ref.so missing"
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    print("| config | MP | bytes | ours e2e ms | ref e2e ms | e2e x | ours device ms | ref device ms | device x | ours MP/s (device) |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for name, data in cases():
        mp, oe, od = time_ours(data, flush)
        re_, rdv = time_ref(data, flush)
        row = dict(config=name, megapixels=round(mp, 2), bytes=len(data), ours_e2e_ms=round(oe, 4),
                   ours_device_ms=round(od, 4), ref_e2e_ms=None if re_ is None else round(re_, 4),
                   ref_device_ms=None if rdv is None else round(rdv, 4))
        rows.append(row)
        fx = "-" if re_ is None else f"{re_ / oe:.2f}"
        fd = "-" if rdv is None else f"{rdv / od:.2f}"
        print(f"| {name} | {mp:.1f} | {len(data)} | {oe:.3f} | {'-' if re_ is None else f'{re_:.3f}'} | {fx} | "
              f"{od:.3f} | {'-' if rdv is None else f'{rdv:.3f}'} | {fd} | {mp / od * 1e3:.0f} |")
    print()
    print(json.dumps({"iters": ITERS, "rows": rows}))


if __name__ == "__main__":
    main()
