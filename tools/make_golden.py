#!/usr/bin/env python
"""Generates tests/golden/reference_digests.json: SHA-256 digests of the quantised-coefficient
planes and the sample planes that the UNMODIFIED reference CUDA decoder
(oracle/_ref/libjpeggpu_ref.so, built from /root/reference/src by oracle/Makefile) produces on
a B200 for the named inputs of tests/inputs.py.  Needs a GPU:

    gpurun -- 'python tools/make_golden.py gpurun_out/reference_digests.json'

The digests pin the CPU oracle in the GPU-less test suite (tests/test_oracle.py).  Each entry
also records the SHA-256 of the input file, so an entry whose input cannot be regenerated
byte-identically (another libjpeg-turbo build) is skipped rather than failed.
"""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

import inputs  # noqa: E402
import oracle_bindings as ob  # noqa: E402
import ref_bindings as rb  # noqa: E402


def digest(arrs):
    h = hashlib.sha256()
    for a in arrs:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden",
                                                                  "reference_digests.json")
    assert rb.available(), "oracle/_ref/libjpeggpu_ref.so missing"
    gold = {}
    for name in inputs.NAMED:
        data = inputs.read_named(name)
        if data is None:
            print("skip", name)
            continue
        # the oracle's parse supplies only the geometry needed to find the reference's
        # coefficient planes inside its scratch buffer; the data is the reference's
        orc = ob.decode(data)
        oi = orc["info"]
        cw = [c.shape[1] for c in orc["planes"]] + [0] * (4 - len(orc["planes"]))
        ch = [c.shape[0] for c in orc["planes"]] + [0] * (4 - len(orc["planes"]))
        comps = [[oi.scans[s].comp_idx[a] for a in range(oi.scans[s].num_components)]
                 for s in range(oi.num_scans)]
        st, rinfo, rplanes, rcoefs = rb.decode_with_reference(
            data, [oi.scans[s].num_segments for s in range(oi.num_scans)], cw, ch, comps)
        assert st == 0, (name, st)
        n = rinfo.num_components
        gold[name] = {
            "input_sha256": hashlib.sha256(data).hexdigest(),
            "input_bytes": len(data),
            "sizes": [[int(rinfo.sizes_x[c]), int(rinfo.sizes_y[c])] for c in range(n)],
            "pixels_sha256": digest(rplanes[:n]),
            "coefficients_sha256": digest(rcoefs[:n]) if rcoefs is not None else None,
        }
        print(name, gold[name]["pixels_sha256"][:12], (gold[name]["coefficients_sha256"] or "-")[:12])
    os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
    with open(out_path, "w") as f:
        json.dump(gold, f, indent=1, sort_keys=True)
    print("wrote", out_path, len(gold), "entries")


if __name__ == "__main__":
    main()
