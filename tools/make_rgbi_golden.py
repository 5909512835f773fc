"""Writes tests/golden/rgbi_digests.json: SHA-256 of the output of the REFERENCE's own
conv_to_rgbi (util/util.h:33-107, compiled from /root/reference into oracle/_ref/libref_util.so by
oracle/Makefile) on the deterministic inputs of tests/test_rgbi_oracle.py.  Run in the build
container (the reference header is not on the GPU box)."""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")]
import oracle_bindings as ob  # noqa: E402
import test_rgbi_oracle as t  # noqa: E402

assert ob.ref_util_available(), "build oracle/_ref first (make -C oracle)"
gold = {"_source": "reference util/util.h conv_to_rgbi via oracle/ref_util_shim.c, gcc -O2, x86-64"}
rc, out = ob.ref_conv_to_rgbi(t.all_triples(), [1, 1, 1], [1, 1, 1])
assert rc == 0
gold["all_triples_444"] = hashlib.sha256(out.tobytes()).hexdigest()
for w, h, fx, fy in t.CASES:
    rc, out = ob.ref_conv_to_rgbi(t.subsampled_case(w, h, fx, fy, seed=w * 131 + h), [fx, 1, 1], [fy, 1, 1])
    assert rc == 0
    gold[f"{w}x{h}_{fx}x{fy}"] = hashlib.sha256(out.tobytes()).hexdigest()
json.dump(gold, open(os.path.join(ROOT, "tests", "golden", "rgbi_digests.json"), "w"), indent=1)
print("wrote", len(gold) - 1, "digests")
