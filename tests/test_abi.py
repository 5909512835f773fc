"""CPU tests of the C-ABI library: it loads without a GPU, exports every symbol the headers
declare, and the host-only entry points (startup, parse_header, get_buffer_size, status strings,
argument checks) behave like the reference's."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import inputs
import jpeggpu_b200 as jb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(jb.LIB_PATH):
        jb.build_library()
    return jb.load_library()


def test_exports_every_declared_symbol(lib):
    declared = set()
    for hdr in ("include/jpeggpu/jpeggpu.h", "include/jpeggpu_b200.h"):
        text = open(os.path.join(ROOT, hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        declared |= set(re.findall(r"\b(jpeggpu_[a-z0-9_]+|is_css_444)\s*\(", text))
    assert declared == set(jb.API_SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name


def test_status_strings(lib):
    want = ["success", "invalid argument", "invalid jpeg", "internal jpeggpu error",
            "jpeg is not supported", "out of host memory", "incomplete bitstream"]
    for i, s in enumerate(want):
        assert lib.jpeggpu_get_status_string(i) == s.encode()


def test_null_arguments(lib):
    assert lib.jpeggpu_decoder_startup(None) == jb.INVALID_ARGUMENT
    assert lib.jpeggpu_set_logging(None, 1) == jb.INVALID_ARGUMENT
    assert lib.jpeggpu_decoder_cleanup(None) == jb.INVALID_ARGUMENT
    assert lib.jpeggpu_decoder_get_buffer_size(None, None) == jb.INVALID_ARGUMENT
    d = jb.Decoder(lib)
    assert lib.jpeggpu_decoder_get_buffer_size(d.handle, None) == jb.INVALID_ARGUMENT
    assert lib.jpeggpu_decoder_decode(d.handle, None, None, 0, None) == jb.INVALID_ARGUMENT
    d.cleanup()


def test_is_css_444(lib):
    s = jb.Subsampling()
    for c in range(4):
        s.x[c] = s.y[c] = 1
    assert lib.is_css_444(s, 3) == 1
    s.x[0] = 2
    assert lib.is_css_444(s, 3) == 0
    assert lib.is_css_444(s, 0) == 0


def test_parse_header_on_cpu(lib):
    data = inputs.pillow_jpeg(w=643, h=477, seed=3, mode="YCbCr", subsampling=2, restart_blocks=4)
    buf = np.frombuffer(data, dtype=np.uint8)
    d = jb.Decoder(lib)
    info = d.parse_header(buf.ctypes.data, len(data), keepalive=buf)
    assert info.num_components == 3
    assert list(info.sizes_x)[:3] == [643, 322, 322]
    assert list(info.sizes_y)[:3] == [477, 239, 239]
    assert list(info.subsampling.x)[:3] == [2, 1, 1]
    assert info.sizes_x[3] == 0 and info.subsampling.x[3] == 0
    assert d.get_buffer_size() > 0
    st = d.scan_stats(0)
    assert st.dus_per_mcu == 6 and st.total_mcus == 41 * 30
    assert st.num_segments == (41 * 30 + 3) // 4
    assert d.transfer_size() < len(data) + 256 * 1024  # + decode LUTs and directories
    w, h = d.coefficient_size()
    assert (w[0], h[0], w[1], h[1]) == (656, 480, 328, 240)
    d.cleanup()


def test_parse_statuses(lib):
    import io
    from PIL import Image
    good = inputs.pillow_jpeg(w=64, h=64, seed=9, mode="YCbCr", subsampling=2)
    d = jb.Decoder(lib)

    def status(b):
        buf = np.frombuffer(b, dtype=np.uint8) if len(b) else np.zeros(1, dtype=np.uint8)
        info = jb.ImgInfo()
        return lib.jpeggpu_decoder_parse_header(d.handle, C.byref(info), buf.ctypes.data, len(b))

    assert status(good) == jb.SUCCESS
    assert status(b"") == jb.INVALID_JPEG
    assert status(good[:-2]) == jb.INVALID_JPEG           # no EOI
    buf = io.BytesIO()
    Image.fromarray(inputs.noise_rgb(64, 64, 1), "RGB").save(buf, "JPEG", progressive=True)
    assert status(buf.getvalue()) == jb.NOT_SUPPORTED
    sof = good.index(b"\xff\xc0")
    assert status(good[:sof + 4] + bytes([12]) + good[sof + 5:]) == jb.NOT_SUPPORTED
    # a failed parse leaves the decoder unusable for size queries until the next good parse
    n = C.c_size_t()
    assert lib.jpeggpu_decoder_get_buffer_size(d.handle, C.byref(n)) == jb.INVALID_ARGUMENT
    assert status(good) == jb.SUCCESS
    assert lib.jpeggpu_decoder_get_buffer_size(d.handle, C.byref(n)) == jb.SUCCESS
    d.cleanup()
