#!/usr/bin/env python
"""Device time of the three ways to reach interleaved RGB for one image (CUDA events, L2 flushed
before every timed call): (a) jpeggpu_decoder_decode to planes only, (b) planes +
jpeggpu_b200_convert_rgbi, (c) fused jpeggpu_b200_decode_rgbi; plus the per-kernel time of the
fused IDCT and the stand-alone converter with their algorithmic bytes and GB/s, and (d) the
reference's route: planes copied to the host + conv_to_rgbi restatement on one core.
usage: rgbi_times.py [file ...]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")]
import torch

import gen_inputs
import jpeggpu_b200 as jb
from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes


def timed(fn, flush, reps=20):
    tot = 0.0
    for _ in range(reps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    return tot / reps * 1e3


def main():
    files = sys.argv[1:] or [gen_inputs.ensure_synth(4032, 3024, 6510), gen_inputs.ensure_img6510()]
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for path in files:
        data = open(path, "rb").read()
        sess = DecodeSession("cuda:0")
        info = sess.parse(pinned_bytes(data))
        planes = sess.alloc_planes()
        h, w = info.sizes_y[0], info.sizes_x[0]
        rgb = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
        sess.transfer()
        ptrs, pitches = [p.data_ptr() for p in planes], [p.stride(0) for p in planes]
        conv = lambda: jb.convert_rgbi(ptrs, pitches, info, rgb.data_ptr(), rgb.stride(0), stream)
        for _ in range(3):
            sess.decode(planes); conv(); sess.decode_rgbi(rgb)
        torch.cuda.synchronize()
        res = {"file": os.path.basename(path), "width": w, "height": h}
        res["planes_us"] = timed(lambda: sess.decode(planes), flush)
        res["planes_then_convert_us"] = timed(lambda: (sess.decode(planes), conv()), flush)
        res["fused_rgbi_us"] = timed(lambda: sess.decode_rgbi(rgb), flush)
        res["convert_only_us"] = timed(conv, flush)
        sess.dec.set_profiling(True)
        acc = {}
        for _ in range(10):
            flush.fill_(1)
            sess.decode_rgbi(rgb)
            torch.cuda.synchronize()
            for k, (ms, n) in sess.dec.kernel_times().items():
                acc[k] = acc.get(k, 0.0) + ms / 10 * 1e3
        sess.dec.set_profiling(False)
        res["fused_kernels_us"] = {k: round(v, 2) for k, v in acc.items()}
        st = sess.dec.scan_stats(0)
        coef_bytes = st.num_data_units * 130
        p_bytes = sum(info.sizes_x[c] * info.sizes_y[c] for c in range(info.num_components))
        fused_bytes = coef_bytes + 3 * w * h
        res["idct_rgbi_bytes"] = fused_bytes
        res["idct_rgbi_gbs"] = fused_bytes / (acc.get("idct_rgbi_kernel", float("nan")) * 1e-6) / 1e9
        res["convert_bytes"] = p_bytes + 3 * w * h
        res["convert_gbs"] = res["convert_bytes"] / (res["convert_only_us"] * 1e-6) / 1e9
        # the reference's route (example/example_tool.c): D2H of the planes + conv_to_rgbi on a core
        try:
            import oracle_bindings as ob
            t0 = time.perf_counter()
            host = [p[:, : info.sizes_x[c]].cpu().numpy() for c, p in enumerate(planes)]
            t1 = time.perf_counter()
            n = info.num_components
            ob.conv_to_rgbi(host, list(info.subsampling.x)[:n], list(info.subsampling.y)[:n])
            t2 = time.perf_counter()
            res["cpu_route_us"] = {"d2h_planes": (t1 - t0) * 1e6, "conv_to_rgbi_1core": (t2 - t1) * 1e6}
        except Exception as e:  # the oracle is optional for this tool
            res["cpu_route_us"] = str(e)
        for k, v in list(res.items()):
            if isinstance(v, float):
                res[k] = round(v, 2)
        print(json.dumps(res))
        sess.close()


if __name__ == "__main__":
    main()
