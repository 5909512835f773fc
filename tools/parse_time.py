#!/usr/bin/env python
"""Host-only timing of parse_header (ours and, if built, the reference's) on pageable and pinned
copies of a file; prints the NUMA layout.  Diagnostic for the e2e host overhead."""
import ctypes as C
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402

import gen_inputs  # noqa: E402
import jpeggpu_b200 as j  # noqa: E402


def timeit(fn, n=300):
    for _ in range(10):
        fn()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    ts.sort()
    return ts[len(ts) // 2] * 1e6, ts[0] * 1e6


def main():
    L = j.load_library()
    try:
        import torch
        have_cuda = torch.cuda.is_available()
    except Exception:
        have_cuda = False
    try:
        import ref_bindings as rb
        ref_ok = rb.available() and have_cuda
    except Exception:
        ref_ok = False
    print("affinity", sorted(os.sched_getaffinity(0)))
    for cmd in (["lscpu"], ["nvidia-smi", "topo", "-m"]):
        try:
            out = subprocess.run(cmd, capture_output=True, text=True, timeout=20).stdout
            for line in out.splitlines():
                if any(k in line for k in ("NUMA", "Model name", "GPU0", "Socket", "Thread", "CPU Affinity")):
                    print("  ", line[:160])
        except Exception as e:
            print(cmd, "failed", e)
    for path in (gen_inputs.ensure_synth(4032, 3024, 6510, 90), gen_inputs.ensure_img6510()):
        data = np.fromfile(path, dtype=np.uint8)
        bufs = {"pageable": data}
        if have_cuda:
            t = torch.empty(data.size, dtype=torch.uint8, pin_memory=True)
            t.numpy()[:] = data
            bufs["pinned"] = t.numpy()
        for kind, arr in bufs.items():
            d = j.Decoder()
            info = j.ImgInfo()
            med, best = timeit(lambda: L.jpeggpu_decoder_parse_header(d.handle, C.byref(info), arr.ctypes.data, arr.size))
            line = f"{os.path.basename(path)} {kind}: ours median {med:.1f} us best {best:.1f} us"
            if ref_ok:
                rd = rb.RefDecoder()
                m2, b2 = timeit(lambda: rd.parse(arr.ctypes.data, arr.size))
                line += f" | reference median {m2:.1f} us best {b2:.1f} us"
                rd.close()
            t0 = time.perf_counter()
            for _ in range(50):
                int(arr.sum(dtype=np.uint64))
            line += f" | numpy sum {(time.perf_counter() - t0) / 50 * 1e6:.1f} us"
            print(line)


if __name__ == "__main__":
    main()
