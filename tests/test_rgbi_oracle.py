"""CPU tests of the colour-conversion oracle (orc_conv_to_rgbi, the restatement of the
reference's util/util.h:33-107) and of the host-side argument checks of the two RGB entry points.
The restatement is PINNED here against the reference's own function, compiled from its header
into oracle/_ref/libref_util.so by oracle/Makefile."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

import inputs
import jpeggpu_b200 as jb
import oracle_bindings as ob

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rgbi_digests.json")
needs_ref = pytest.mark.skipif(not ob.ref_util_available(), reason="oracle/_ref/libref_util.so not built")


def all_triples():
    """4096x4096 4:4:4 planes holding every (y, cb, cr) once."""
    v = np.arange(1 << 24, dtype=np.uint32)
    mk = lambda s: ((v >> s) & 255).astype(np.uint8).reshape(4096, 4096)
    return [mk(0), mk(8), mk(16)]


def subsampled_case(w, h, fx, fy, seed):
    rng = np.random.RandomState(seed)
    cw, ch = (w + fx - 1) // fx, (h + fy - 1) // fy
    return [rng.randint(0, 256, (h, w), dtype=np.uint8), rng.randint(0, 256, (ch, cw), dtype=np.uint8),
            rng.randint(0, 256, (ch, cw), dtype=np.uint8)]


CASES = [(64, 48, 2, 2), (17, 13, 2, 2), (250, 3, 2, 1), (33, 31, 1, 2), (1, 1, 2, 2), (40, 9, 4, 1),
         (29, 7, 1, 1)]


@needs_ref
def test_oracle_equals_reference_function_on_every_triple():
    planes = all_triples()
    rc_a, a = ob.conv_to_rgbi(planes, [1, 1, 1], [1, 1, 1])
    rc_b, b = ob.ref_conv_to_rgbi(planes, [1, 1, 1], [1, 1, 1])
    assert rc_a == 0 and rc_b == 0
    assert np.array_equal(a, b)


@needs_ref
@pytest.mark.parametrize("w,h,fx,fy", CASES)
def test_oracle_equals_reference_function_subsampled(w, h, fx, fy):
    planes = subsampled_case(w, h, fx, fy, seed=w * 131 + h)
    rc_a, a = ob.conv_to_rgbi(planes, [fx, 1, 1], [fy, 1, 1])
    rc_b, b = ob.ref_conv_to_rgbi(planes, [fx, 1, 1], [fy, 1, 1])
    assert rc_a == 0 and rc_b == 0
    assert np.array_equal(a, b)


@needs_ref
def test_oracle_failure_codes_follow_the_reference():
    p = subsampled_case(16, 16, 1, 1, 3)
    grey = [p[0]]
    for n, fx, fy in [(3, [1, 2, 2], [1, 2, 2]), (3, [2, 1, 2], [2, 1, 1]), (3, [2, 2, 2], [2, 2, 2])]:
        a = ob.conv_to_rgbi(p, fx, fy)[0]
        b = ob.ref_conv_to_rgbi(p, fx, fy)[0]
        assert (a != 0) == (b != 0), (fx, fy)
    assert ob.conv_to_rgbi(grey, [1], [1])[0] == 0
    rc, out = ob.conv_to_rgbi(grey, [1], [1])
    assert np.array_equal(out, np.repeat(grey[0][:, :, None], 3, axis=2))


def test_oracle_against_committed_digests():
    """Digests written by tools/make_rgbi_golden.py from the REFERENCE function in the build
    container; this check needs no reference at run time."""
    gold = json.load(open(GOLDEN))
    rc, a = ob.conv_to_rgbi(all_triples(), [1, 1, 1], [1, 1, 1])
    assert rc == 0 and hashlib.sha256(a.tobytes()).hexdigest() == gold["all_triples_444"]
    for w, h, fx, fy in CASES:
        rc, a = ob.conv_to_rgbi(subsampled_case(w, h, fx, fy, seed=w * 131 + h), [fx, 1, 1], [fy, 1, 1])
        assert rc == 0
        assert hashlib.sha256(a.tobytes()).hexdigest() == gold[f"{w}x{h}_{fx}x{fy}"]


def test_float_model_matches_numpy_float32():
    """Independent statement of the same arithmetic in numpy float32 (no FMA)."""
    planes = subsampled_case(96, 64, 2, 2, 5)
    rc, a = ob.conv_to_rgbi(planes, [2, 1, 1], [2, 1, 1])
    y = planes[0].astype(np.float32)
    cb = np.repeat(np.repeat(planes[1], 2, 0), 2, 1).astype(np.float32) - np.float32(128)
    cr = np.repeat(np.repeat(planes[2], 2, 0), 2, 1).astype(np.float32) - np.float32(128)
    f = np.float32
    r = y + f(1.402) * cr
    g = (y - f(.344136) * cb) - f(.714136) * cr
    b = y + f(1.772) * cb
    rnd = lambda v: np.clip(np.where(v >= 0, np.floor(v + f(0.5)), -np.floor(-v + f(0.5))), 0, 255).astype(np.uint8)
    assert np.array_equal(a, np.stack([rnd(r), rnd(g), rnd(b)], axis=2))


# ---- host-side checks of the product entry points (no GPU work is enqueued) ---------------------

@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(jb.LIB_PATH):
        jb.build_library()
    return jb.load_library()


def _parsed(lib, data):
    dec = jb.Decoder(lib)
    buf = np.frombuffer(data, dtype=np.uint8).copy()
    info = dec.parse_header(buf.ctypes.data, buf.size, keepalive=buf)
    return dec, info


def test_decode_rgbi_argument_checks(lib):
    dec, info = _parsed(lib, inputs.pillow_jpeg(w=64, h=48, seed=1, subsampling=2))
    h = dec.handle
    assert lib.jpeggpu_b200_decode_rgbi(None, 1, 192, 1, 1, None) == jb.INVALID_ARGUMENT
    assert lib.jpeggpu_b200_decode_rgbi(h, None, 192, 1, 1, None) == jb.INVALID_ARGUMENT
    assert lib.jpeggpu_b200_decode_rgbi(h, 1, 3 * 64 - 1, 1, 1, None) == jb.INVALID_ARGUMENT
    fresh = jb.Decoder(lib)
    assert lib.jpeggpu_b200_decode_rgbi(fresh.handle, 1, 192, 1, 1, None) == jb.INVALID_ARGUMENT


def test_decode_rgbi_refuses_layouts_outside_conv_to_rgbi(lib):
    # four components (util/util.h:43-46)
    dec, _ = _parsed(lib, inputs.pillow_jpeg(w=64, h=48, seed=2, mode="CMYK"))
    assert lib.jpeggpu_b200_decode_rgbi(dec.handle, 1, 1024, 1, 1, None) == jb.NOT_SUPPORTED
    # three components in three scans: planes first, then jpeggpu_b200_convert_rgbi
    import jpeg_transcode as jt
    import oracle_bindings
    src = inputs.pillow_jpeg(w=64, h=48, seed=3, subsampling=2)
    ref = oracle_bindings.decode(src)
    multi = jt.transcode(src, ref["planes"], scans=[[0], [1], [2]])
    dec, _ = _parsed(lib, multi)
    assert lib.jpeggpu_b200_decode_rgbi(dec.handle, 1, 1024, 1, 1, None) == jb.NOT_SUPPORTED


def test_convert_rgbi_argument_checks(lib):
    img, info = jb.Img(), jb.ImgInfo()
    info.num_components = 3
    for c in range(3):
        info.subsampling.x[c] = info.subsampling.y[c] = 1
        info.sizes_x[c], info.sizes_y[c] = 16, 16
        img.image[c] = 4096
        img.pitch[c] = 16
    f = lib.jpeggpu_b200_convert_rgbi
    assert f(None, C.byref(info), 1, 48, None) == jb.INVALID_ARGUMENT
    assert f(C.byref(img), None, 1, 48, None) == jb.INVALID_ARGUMENT
    assert f(C.byref(img), C.byref(info), None, 48, None) == jb.INVALID_ARGUMENT
    assert f(C.byref(img), C.byref(info), 1, 47, None) == jb.INVALID_ARGUMENT
    info.num_components = 4
    assert f(C.byref(img), C.byref(info), 1, 48, None) == jb.NOT_SUPPORTED
    info.num_components = 2
    assert f(C.byref(img), C.byref(info), 1, 48, None) == jb.NOT_SUPPORTED
    info.num_components = 3
    info.subsampling.x[1] = 2  # unequal / non-1x1 chroma
    assert f(C.byref(img), C.byref(info), 1, 48, None) == jb.NOT_SUPPORTED
    info.subsampling.x[1] = 1
    img.pitch[2] = 15
    assert f(C.byref(img), C.byref(info), 1, 48, None) == jb.INVALID_ARGUMENT
