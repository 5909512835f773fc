"""CPU: parse_header (headers + marker walk, csrc/parser.cpp) on malformed files returns a status
and never crashes, over-reads or loops; statuses stay inside the reference's set
(include/jpeggpu/jpeggpu.h:38-52).  Mutations: the entropy-data ones of test_gpu_robustness plus
byte flips, truncations and length-field damage in the marker segments."""
import os

import numpy as np
import pytest

import inputs
import jpeggpu_b200 as jb
import test_gpu_robustness as rb


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(jb.LIB_PATH):
        jb.build_library()
    return jb.load_library()


def parse_status(lib, dec, data, defer):
    # the buffer is exactly as long as the file and sits at the END of a page-sized allocation's
    # logical extent (numpy copy): an over-read of more than the documented 1 byte would at least
    # touch foreign bytes; statuses are what we assert on
    buf = np.frombuffer(data, dtype=np.uint8).copy()
    lib.jpeggpu_b200_set_deferred_walk(dec.handle, int(defer))
    info = jb.ImgInfo()
    import ctypes as C
    st = lib.jpeggpu_decoder_parse_header(dec.handle, C.byref(info), buf.ctypes.data, buf.size)
    assert 0 <= st <= jb.INCOMPLETE_BITSTREAM
    if st == jb.SUCCESS:
        n = C.c_size_t()
        assert lib.jpeggpu_decoder_get_buffer_size(dec.handle, C.byref(n)) == jb.SUCCESS
        assert 0 < n.value < (1 << 33)
        assert 1 <= info.num_components <= 4
        for c in range(info.num_components):
            assert 0 < info.sizes_x[c] <= 65535 and 0 < info.sizes_y[c] <= 65535
    return st


@pytest.mark.parametrize("src", list(rb.SOURCES))
@pytest.mark.parametrize("defer", [0, 1])
def test_entropy_mutations(lib, src, defer):
    good = inputs.pillow_jpeg(**rb.SOURCES[src])
    dec = jb.Decoder(lib)
    for kind in rb.KINDS:
        for seed in range(3):
            parse_status(lib, dec, rb.mutate(good, kind, seed), defer)
    assert parse_status(lib, dec, good, defer) == jb.SUCCESS
    dec.cleanup()


@pytest.mark.parametrize("src", list(rb.SOURCES))
def test_header_mutations(lib, src):
    good = inputs.pillow_jpeg(**rb.SOURCES[src])
    lo, _ = rb.scan_range(good)
    dec = jb.Decoder(lib)
    rng = np.random.RandomState(7)
    seen = set()
    for trial in range(400):
        b = bytearray(good)
        for _ in range(rng.randint(1, 4)):
            i = rng.randint(2, lo)
            b[i] = rng.randint(256) if rng.rand() < 0.5 else b[i] ^ (1 << rng.randint(8))
        seen.add(parse_status(lib, dec, bytes(b), trial & 1))
    for cut in list(range(0, lo + 8)) + [lo + 100, len(good) - 2, len(good) - 1]:
        seen.add(parse_status(lib, dec, good[:cut], cut & 1))
    assert jb.SUCCESS in seen and len(seen) >= 3  # the mutations do reach the error paths
    assert parse_status(lib, dec, good, 1) == jb.SUCCESS
    dec.cleanup()


def test_degenerate_buffers(lib):
    dec = jb.Decoder(lib)
    for data in (b"\xff", b"\xff\xd8", b"\xff\xd8\xff", b"\xff\xd8\xff\xd9", b"\x00" * 64, b"\xff" * 64,
                 b"\xff\xd8" + b"\xff" * 100, b"\xff\xd8\xff\xda\x00\x02", b"\xff\xd8\xff\xc0\x00\x0b\x08"):
        assert parse_status(lib, dec, data, 1) != jb.SUCCESS
    dec.cleanup()
