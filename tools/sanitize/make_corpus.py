"""Writes the mutated-JPEG corpus for tools/sanitize_parser.sh into the directory given as argv[1]:
the entropy-data mutations of tests/test_gpu_robustness.py, random header bytes, every prefix."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")]
import numpy as np  # noqa: E402

import inputs  # noqa: E402
import test_gpu_robustness as rb  # noqa: E402

out = sys.argv[1]
os.makedirs(out, exist_ok=True)
n = 0
srcs = dict(rb.SOURCES)
srcs["big"] = dict(w=2048, h=1536, seed=64, mode="YCbCr", subsampling=0, quality=97)  # > 1 MiB scan
for name, kw in srcs.items():
    good = inputs.pillow_jpeg(**kw)
    lo, hi = rb.scan_range(good)
    files = {f"{name}_good": good}
    for kind in rb.KINDS:
        for seed in range(2):
            files[f"{name}_{kind}_{seed}"] = rb.mutate(good, kind, seed)
    rng = np.random.RandomState(1)
    for t in range(60):
        b = bytearray(good)
        for _ in range(rng.randint(1, 4)):
            b[rng.randint(2, lo)] = rng.randint(256)
        files[f"{name}_hdr_{t}"] = bytes(b)
    for cut in list(range(0, lo + 8, 3)) + [lo + 100, len(good) - 2, len(good) - 1]:
        files[f"{name}_cut_{cut}"] = good[:cut]
    for k, v in files.items():
        open(os.path.join(out, k + ".jpg"), "wb").write(v)
        n += 1
print(n, "files in", out)
