"""CPU tests of the product's host-side logic and of the code the kernels share with the host
(parser, scan planner, LUT builder, Huffman sync/verify/write thread bodies), run sequentially by
tests/host_emul/emul.cpp and compared with the oracle bit-exactly."""
import numpy as np
import pytest

import emul_bindings as eb
import inputs
import oracle_bindings as ob

CASES = [
    dict(w=640, h=480, seed=1, mode="YCbCr", subsampling=0),
    dict(w=640, h=480, seed=2, mode="YCbCr", subsampling=1),
    dict(w=643, h=477, seed=3, mode="YCbCr", subsampling=2),
    dict(w=640, h=480, seed=4, mode="L"),
    dict(w=320, h=240, seed=5, mode="CMYK"),
    dict(w=500, h=375, seed=6, mode="YCbCr", subsampling=2, restart_blocks=1),
    dict(w=500, h=375, seed=7, mode="YCbCr", subsampling=2, restart_blocks=7),
    dict(w=500, h=375, seed=8, mode="YCbCr", subsampling=0, optimize=True, quality=98),
    dict(w=500, h=375, seed=9, mode="YCbCr", subsampling=2, optimize=True, quality=20),
    dict(w=8, h=8, seed=10, mode="L"),
    dict(w=1, h=1, seed=11, mode="YCbCr", subsampling=2),
    dict(w=1200, h=900, seed=12, mode="YCbCr", subsampling=2, quality=100),
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items()))
def test_emulated_pipeline_matches_oracle(case):
    data = inputs.pillow_jpeg(**case)
    ref = ob.decode(data)
    assert ref["status"] == 0
    for s in range(ref["info"].num_scans):
        st, coef, stats, _ = eb.decode_scan(data, s, ref["stream"][s].size)
        assert st == 0
        assert stats[4] == 0 or True  # records may still be dirty before the fix-up pass
        assert np.array_equal(coef, ref["stream"][s]), f"scan {s}: {(coef != ref['stream'][s]).sum()} differ"


def test_emulated_destuff_matches_oracle():
    data = inputs.pillow_jpeg(w=500, h=375, seed=7, mode="YCbCr", subsampling=2, restart_blocks=3,
                              quality=100)
    st, info = ob.parse(data)
    sc = info.scans[0]
    want, off, ln = ob.destuff(data[sc.begin:sc.end], 64)
    st, coef, stats, dst = eb.decode_scan(data, 0, sc.mcus_x * sc.mcus_y * sc.dus_per_mcu * 64,
                                          destuffed_bytes=len(want))
    assert st == 0
    assert stats[3] * 64 == len(want)
    assert np.array_equal(dst, want)


def test_parser_agrees_with_oracle_on_geometry():
    for case in CASES:
        data = inputs.pillow_jpeg(**case)
        st, out = eb.parse(data)
        ost, info = ob.parse(data)
        assert st == ost == 0
        assert (out[0], out[1], out[2], out[3], out[4]) == (
            info.width, info.height, info.num_components, info.num_scans, info.restart_interval)
        assert out[5] == info.scans[0].num_segments
        assert (out[8], out[9]) == (info.scans[0].begin, info.scans[0].end)


def test_parser_statuses_match_oracle():
    import io
    from PIL import Image
    good = inputs.pillow_jpeg(w=64, h=64, seed=9, mode="YCbCr", subsampling=2)
    samples = [b"", b"\xff\xd8", b"\x00" * 16, good[:50], good[:300], good[:-2], good + b"trailer"]
    buf = io.BytesIO()
    Image.fromarray(inputs.noise_rgb(64, 64, 1), "RGB").save(buf, "JPEG", progressive=True)
    samples.append(buf.getvalue())
    # 12-bit precision and 16-bit DQT are NOT_SUPPORTED
    sof = good.index(b"\xff\xc0")
    samples.append(good[:sof + 4] + bytes([12]) + good[sof + 5:])
    dqt = good.index(b"\xff\xdb")
    samples.append(good[:dqt + 4] + bytes([0x10]) + good[dqt + 5:])
    for s in samples:
        st, _ = eb.parse(s) if len(s) else (2, None)
        ost, _ = ob.parse(s)
        assert st == ost, (len(s), st, ost)


@pytest.mark.parametrize("case", [
    dict(w=643, h=477, seed=3, mode="YCbCr", subsampling=2),
    dict(w=500, h=375, seed=7, mode="YCbCr", subsampling=2, restart_blocks=7),
    dict(w=500, h=375, seed=8, mode="YCbCr", subsampling=0, optimize=True, quality=98),
    dict(w=500, h=375, seed=9, mode="YCbCr", subsampling=2, optimize=True, quality=20),
    dict(w=320, h=240, seed=5, mode="CMYK"),
], ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items()))
def test_pair_steps_equal_single_steps(case):
    """huff_sync consumes two symbols per table probe inside its fast zone (zig-zag index < 48,
    more than 10 bits to go); records must equal single stepping for every subsequence from the
    cold state and from arbitrary (speculative) entry states."""
    data = inputs.pillow_jpeg(**case)
    bad, pairs = eb.pair_equivalence(data, tries=4)
    assert pairs > 0, "the tables hold no pair entries: nothing was tested"
    assert bad == 0


def test_img6510_emulated():
    data = inputs.img6510()
    if data is None:
        pytest.skip("IMG_6510.JPG not staged")
    ref = ob.decode(data)
    st, coef, stats, _ = eb.decode_scan(data, 0, ref["stream"][0].size)
    assert st == 0
    assert np.array_equal(coef, ref["stream"][0])
    st, out = eb.parse(data)
    assert out[5] == 189 and out[7] == 2907282


@pytest.mark.parametrize("case", [
    dict(w=1200, h=900, seed=21, mode="YCbCr", subsampling=2),
    dict(w=1200, h=900, seed=22, mode="YCbCr", subsampling=2, restart_blocks=75),
    dict(w=800, h=600, seed=23, mode="L"),
], ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items()))
def test_link_repairs_damaged_records(case):
    """huff_link's fallback (relink_from + re-verification): the emulator damages sync records
    after the sync pass; the repaired decode must still equal the oracle and must have gone
    through the slow path."""
    data = inputs.pillow_jpeg(**case)
    ref = ob.decode(data)
    assert ref["status"] == 0
    st, coef, stats, _ = eb.decode_scan(data, 0, ref["stream"][0].size, corrupt_records=True)
    assert st == 0
    assert stats[2] > 0, "the repair path did not run"
    assert np.array_equal(coef, ref["stream"][0])


@pytest.mark.parametrize("defs", ["-DJB2_WRITE_PARTS=4", "-DJB2_SUB_BYTES=128 -DJB2_OVERLAP_SUBS=4"])
def test_emulated_pipeline_tuning_variants(defs):
    """The compile-time tunables (threads per subsequence in the write pass, subsequence size,
    overlap) keep the pipeline bit-exact: emulator rebuilt with the definitions in a subprocess."""
    import os
    import subprocess
    import sys
    code = (
        "import sys, numpy as np\n"
        "import emul_bindings as eb, inputs, oracle_bindings as ob\n"
        "for kw in (dict(w=643, h=477, seed=3, mode='YCbCr', subsampling=2),\n"
        "           dict(w=500, h=375, seed=7, mode='YCbCr', subsampling=2, restart_blocks=7),\n"
        "           dict(w=320, h=240, seed=5, mode='CMYK')):\n"
        "    data = inputs.pillow_jpeg(**kw)\n"
        "    ref = ob.decode(data)\n"
        "    for s in range(ref['info'].num_scans):\n"
        "        st, coef, stats, _ = eb.decode_scan(data, s, ref['stream'][s].size)\n"
        "        assert st == 0 and np.array_equal(coef, ref['stream'][s]), (kw, s)\n"
        "print('ok')\n")
    here = os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, JB2_DEFS=defs,
               PYTHONPATH=os.pathsep.join([here, os.path.join(os.path.dirname(here), "tools"),
                                           os.path.dirname(here)]))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0 and "ok" in r.stdout, r.stdout[-1500:] + r.stderr[-1500:]


# ---- deferred / parallel marker walk (csrc/parser.cpp) ------------------------------------------

def _big_scan_file(seed, w=2048, h=1536, quality=97):
    """> 1 MiB of entropy-coded data in one scan without DRI (the parallel walk's precondition)."""
    data = inputs.pillow_jpeg(w=w, h=h, seed=seed, mode="YCbCr", subsampling=0, quality=quality)
    assert len(data) > (1 << 20) + (1 << 17), len(data)
    return data


@pytest.mark.parametrize("threads", [False, True])
def test_deferred_walk_equals_eager_walk(threads):
    for seed in (61, 62):
        data = _big_scan_file(seed)
        for _ in range(3):  # the pool is reused across images
            rc, used, st = eb.walk_compare(data, threads)
            assert (rc, st) == (0, 0)
            assert used == (2 if threads else 0)  # 2: the stitched parallel result was kept
    small = inputs.pillow_jpeg(w=640, h=480, seed=63, mode="YCbCr", subsampling=2)
    rc, used, st = eb.walk_compare(small, threads)
    assert (rc, st, used) == (0, 0, 0)


def test_parallel_walk_falls_back_on_markers_and_damage():
    """Anything but data and stuffed pairs inside a counting block (restart markers without DRI,
    fill bytes, an early EOI, truncation) must give exactly what the sequential walk gives."""
    import test_gpu_robustness as rb
    good = _big_scan_file(64)
    lo, hi = rb.scan_range(good)
    variants = [rb.mutate(good, kind, seed) for kind in rb.KINDS for seed in range(2)]
    b = bytearray(good)
    b[lo + 300000: lo + 300002] = b"\xff\xd3"  # one stray RSTn deep inside a counting block
    variants.append(bytes(b))
    b = bytearray(good)
    cut = lo + (4096 * 37) - 1                 # a stuffed pair straddling a tile boundary
    b[cut: cut + 2] = b"\xff\x00"
    variants.append(bytes(b))
    b = bytearray(good)
    cut = lo + 65536 * 5 - 1                   # ... and a block boundary
    b[cut: cut + 2] = b"\xff\x00"
    variants.append(bytes(b))
    kept = 0
    for v in variants:
        rc, used, st = eb.walk_compare(v, True)
        assert rc == 0, (rc, st)
        kept += used == 2
    assert kept >= 2  # the straddling stuffed pairs are ordinary data: parallel result kept
