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


def _stats_tuple(st):
    return (st.num_components, st.num_segments, st.dus_per_mcu, st.total_mcus, st.scan_bytes,
            st.destuffed_bytes, st.num_subsequences, st.num_data_units)


@pytest.mark.parametrize("mode", [1, 3], ids=["deferred_host_walk", "device_walk"])
@pytest.mark.parametrize("tail", [b"", b"\x00" * 7], ids=["ends_with_eoi", "trailing_bytes"])
def test_deferred_walk_plans_from_bounds(lib, tail, mode):
    """include/jpeggpu_b200.h jpeggpu_b200_set_deferred_walk: a single-scan file is planned from
    upper bounds at parse time; the exact counts (after the host walk that a statistics call - or,
    in mode 1, transfer - triggers) equal those of the eager parse.  Mode 1 (host walk behind the
    upload) needs a file without DRI that ends with EOI: one with bytes behind EOI is parsed
    eagerly.  Mode 3 (the device finds the end of the scan, the default) takes both."""
    data = inputs.pillow_jpeg(w=1600, h=1200, seed=5, mode="YCbCr", subsampling=2) + tail
    assert len(data) > 96 * 1024
    buf = np.frombuffer(data, dtype=np.uint8)
    eager, lazy = jb.Decoder(lib), jb.Decoder(lib)
    eager.set_deferred_walk(False)
    lazy.set_deferred_walk(mode)
    ie = eager.parse_header(buf.ctypes.data, len(data), keepalive=buf)
    il = lazy.parse_header(buf.ctypes.data, len(data), keepalive=buf)
    assert bytes(ie) == bytes(il)
    exact, bound = eager.get_buffer_size(), lazy.get_buffer_size()
    if tail and mode == 1:
        assert bound == exact
    else:
        assert exact <= bound <= exact * 1.01
    assert _stats_tuple(lazy.scan_stats(0)) == _stats_tuple(eager.scan_stats(0))
    assert abs(lazy.transfer_size() - eager.transfer_size()) <= 8192  # capacity slack of the planned tables
    assert lazy.get_buffer_size() == bound  # the plan does not move after the walk
    eager.cleanup()
    lazy.cleanup()


def test_deferred_walk_reports_late_errors(lib):
    """A marker in the middle of the entropy data of an otherwise eligible file: the eager parse
    sees what follows the scan at parse time, the deferred one when the walk runs."""
    data = bytearray(inputs.pillow_jpeg(w=1600, h=1200, seed=6, mode="YCbCr", subsampling=2))
    mid = len(data) // 2
    data[mid:mid + 2] = b"\xff\xc2"  # SOF2 where entropy data should be
    buf = np.frombuffer(bytes(data), dtype=np.uint8)
    eager, lazy = jb.Decoder(lib), jb.Decoder(lib)
    eager.set_deferred_walk(False)
    info = jb.ImgInfo()
    st_e = lib.jpeggpu_decoder_parse_header(eager.handle, C.byref(info), buf.ctypes.data, len(buf))
    assert st_e != jb.SUCCESS
    st_l = lib.jpeggpu_decoder_parse_header(lazy.handle, C.byref(info), buf.ctypes.data, len(buf))
    assert st_l == jb.SUCCESS
    stats = jb.ScanStats()
    assert lib.jpeggpu_b200_get_scan_stats(lazy.handle, 0, C.byref(stats), None) == st_e
    n = C.c_size_t()
    assert lib.jpeggpu_decoder_get_buffer_size(lazy.handle, C.byref(n)) == jb.INVALID_ARGUMENT
    eager.cleanup()
    lazy.cleanup()


def test_destuff_static_rejects_bad_arguments_without_touching_the_device(capfd):
    """jpeggpu_b200_destuff_static checks its arguments before any CUDA call: NULL pointers, sizes of
    zero, pointers that are not 16-byte aligned and scans beyond the static scratch's maximum are
    JPEGGPU_INVALID_ARGUMENT (the last one with a line on stderr that names the limit)."""
    import jpeggpu_b200 as jb
    lib = jb.load_library()
    call = lib.jpeggpu_b200_destuff_static
    a = 0x10000  # never dereferenced
    assert call(None, 4096, a, 64, a, 16, a, None) == jb.INVALID_ARGUMENT
    assert call(a, 4096, None, 64, a, 16, a, None) == jb.INVALID_ARGUMENT
    assert call(a, 4096, a, 64, None, 16, a, None) == jb.INVALID_ARGUMENT
    assert call(a, 4096, a, 64, a, 16, None, None) == jb.INVALID_ARGUMENT
    assert call(a, 0, a, 64, a, 16, a, None) == jb.INVALID_ARGUMENT
    assert call(a, 4096, a, 0, a, 16, a, None) == jb.INVALID_ARGUMENT
    assert call(a, 4096, a, 64, a, 0, a, None) == jb.INVALID_ARGUMENT
    assert call(a + 8, 4096, a, 64, a, 16, a, None) == jb.INVALID_ARGUMENT
    assert call(a, 4096, a + 4, 64, a, 16, a, None) == jb.INVALID_ARGUMENT
    capfd.readouterr()
    assert call(a, jb.DESTUFF_STATIC_MAX_BYTES + 1, a, 64, a, 16, a, None) == jb.INVALID_ARGUMENT
    assert "JPEGGPU_B200_DESTUFF_STATIC_MAX_BYTES" in capfd.readouterr().err
