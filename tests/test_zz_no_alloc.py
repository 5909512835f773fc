"""jpeggpu_b200_destuff_static (include/jpeggpu_b200.h): the destuff kernel with its scratch in a
static device array - an entry point that allocates nothing, for callers that cannot call
cudaMalloc.  It has no counterpart in the reference;
its outputs must equal, bit for bit, what the existing entry point (jpeggpu_decoder_decode) leaves
in its scratch buffer for the same scan: destuffed words, restart-segment directory, scan totals.
Called several times in a row on purpose: every call shares the one static array, which keeps the
previous call's contents.  This file is collected last."""
import ctypes as C

import numpy as np
import pytest

import inputs
import oracle_bindings as ob


def _existing_path(data: bytes):
    """Decode through the reference-shaped API; returns what its destuff kernel wrote."""
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    sess = DecodeSession("cuda:0")
    sess.parse(pinned_bytes(data))
    planes = sess.alloc_planes()
    sess.transfer()
    sess.decode(planes)
    torch.cuda.synchronize()
    st = sess.dec.scan_stats(0)
    info = sess.read_scratch(st.off_scan_info, 16).view(np.uint32).copy()
    n_subs, n_segs = int(info[0]), int(info[1])
    words = sess.read_scratch(st.off_destuffed, n_subs * 64).view(np.uint32).copy()
    segs = sess.read_scratch(st.off_segments, 16 * n_segs).view(np.uint32).reshape(-1, 4).copy()
    sess.close()
    return info, words, segs


def _static_path(lib, data: bytes, scan_begin: int, slack_subs=64, slack_segs=16):
    """The new entry point on a device copy of the bytes behind the SOS header.  Every buffer is a
    torch tensor made BEFORE the call; the call itself gets raw pointers and sizes only."""
    import torch
    import jpeggpu_b200 as jb
    scan = np.frombuffer(data, dtype=np.uint8)[scan_begin:]
    d_scan = torch.from_numpy(scan.copy()).to("cuda:0")
    max_subs = (len(scan) + 63) // 64 + slack_subs
    # upper bound on restart segments: one per RSTn marker + the last one
    ff = np.flatnonzero(scan[:-1] == 0xff)
    max_segs = int(np.count_nonzero((scan[ff + 1] >= 0xd0) & (scan[ff + 1] <= 0xd7))) + slack_segs
    max_subs += max_segs
    d_out = torch.full((max_subs * 16,), 0x5a5a5a5a, dtype=torch.int32, device="cuda:0")
    d_segs = torch.full((max_segs, 4), -1, dtype=torch.int32, device="cuda:0")
    d_info = torch.full((4,), -1, dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    st = lib.jpeggpu_b200_destuff_static(d_scan.data_ptr(), d_scan.numel(), d_out.data_ptr(), max_subs,
                                         d_segs.data_ptr(), max_segs, d_info.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream)
    assert st == jb.SUCCESS, st
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    info = d_info.cpu().numpy().view(np.uint32)
    n_subs, n_segs = int(info[0]), int(info[1])
    words = d_out.cpu().numpy().view(np.uint32)
    assert np.all(words[n_subs * 16:] == 0x5a5a5a5a), "wrote behind the subsequences it reported"
    segs = d_segs.cpu().numpy().view(np.uint32)
    return info, words[:n_subs * 16], segs[:n_segs]


def _compare(lib, data: bytes, label: str):
    st, oinfo = ob.parse(data)
    assert st == 0
    begin = oinfo.scans[0].begin
    want_info, want_words, want_segs = _existing_path(data)
    got_info, got_words, got_segs = _static_path(lib, data, begin)
    assert np.array_equal(got_info, want_info), (label, got_info, want_info)
    assert np.array_equal(got_words, want_words), f"{label}: {(got_words != want_words).sum()} words differ"
    assert np.array_equal(got_segs[:, :2], want_segs[:, :2]), label
    return int(got_info[0])


@pytest.mark.gpu
def test_static_scratch_destuff_matches_the_decode_path_call_after_call():
    import jpeggpu_b200 as jb
    lib = jb.load_library()
    big = inputs.pillow_jpeg(w=1600, h=1200, seed=91, mode="YCbCr", subsampling=2, quality=92, restart_blocks=100)
    small = inputs.pillow_jpeg(w=1024, h=768, seed=92, mode="YCbCr", subsampling=2, quality=90)
    dense = inputs.pillow_jpeg(w=1600, h=1200, seed=93, mode="YCbCr", subsampling=2, quality=90, restart_blocks=1)
    # two calls in a row and more: large -> small (fewer tiles than the previous call left
    # descriptors for) -> the first one again -> many closings per tile
    n1 = _compare(lib, big, "DRI 100, first call")
    n2 = _compare(lib, small, "no DRI, second call")
    n3 = _compare(lib, big, "DRI 100, third call")
    _compare(lib, dense, "DRI 1, fourth call")
    assert n1 == n3 and n2 < n1


@pytest.mark.gpu
def test_static_scratch_destuff_rejects_what_it_cannot_take(capfd):
    import torch
    import jpeggpu_b200 as jb
    lib = jb.load_library()
    buf = torch.zeros(4096, dtype=torch.uint8, device="cuda:0")
    out = torch.zeros(4096, dtype=torch.int32, device="cuda:0")
    p, o = buf.data_ptr(), out.data_ptr()
    call = lib.jpeggpu_b200_destuff_static
    # beyond the maximum the static array is sized for: refused before anything is enqueued
    assert call(p, jb.DESTUFF_STATIC_MAX_BYTES + 1, o, 64, o + 8192, 16, o + 4096, None) == jb.INVALID_ARGUMENT
    assert "JPEGGPU_B200_DESTUFF_STATIC_MAX_BYTES" in capfd.readouterr().err
    assert call(None, 4096, o, 64, o + 8192, 16, o + 4096, None) == jb.INVALID_ARGUMENT
    assert call(p, 0, o, 64, o + 8192, 16, o + 4096, None) == jb.INVALID_ARGUMENT
    assert call(p + 1, 4000, o, 64, o + 8192, 16, o + 4096, None) == jb.INVALID_ARGUMENT   # alignment
    assert call(p, 4096, o, 0, o + 8192, 16, o + 4096, None) == jb.INVALID_ARGUMENT
    # a buffer of zeros is a scan without a terminating marker: 4096 data bytes, flagged
    assert call(p, 4096, o, 64, o + 8192, 16, o + 4096, None) == jb.SUCCESS
    torch.cuda.synchronize()
    info = out[1024:1028].cpu().numpy().view(np.uint32)
    assert list(info) == [64, 1, 4096, 1]  # 64 subsequences, one segment, JPEGGPU_B200_SCAN_UNTERMINATED
