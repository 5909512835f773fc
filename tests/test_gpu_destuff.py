"""Intermediate artefacts of the CUDA path against the CPU oracle (runs on the B200 box, -m gpu):
  * the destuffed scan + restart-segment directory the destuff kernel builds with its on-device
    look-back (reference: src/decode_destuff.cu:152-361 fed by the host walk src/reader.cpp:443-489),
    compared byte for byte with the oracle's destuffer on DRI / no-DRI / fill-byte / marker-polluted
    / truncated inputs, in every walk mode;
  * the Huffman sync records: every subsequence must start where its predecessor stopped, and the
    control block must show that no record needed the sequential repair (a sync regression that
    falls into it stays bit-exact but is many times slower)."""
import numpy as np
import pytest

import inputs
import oracle_bindings as ob

pytestmark = pytest.mark.gpu

CTRL_NAMES = ["relinked", "slow_repairs", "link_rounds", "link_verify_ns", "need_verify"]


def with_fill_bytes(data: bytes):
    """Insert 0xff fill bytes in front of every restart marker (T.81 B.1.1.2 allows them; the
    reference's walk - and therefore the oracle's - takes 0xff 0xff for the end of the scan,
    src/reader.cpp:462-472, so the expected output comes from the file without them).
    Returns (file, bytes inserted)."""
    st, info = ob.parse(data)
    sc = info.scans[0]
    body = bytearray()
    scan = data[sc.begin:sc.end]
    i = 0
    k = 0
    while i < len(scan):
        if scan[i] == 0xff and i + 1 < len(scan) and 0xd0 <= scan[i + 1] <= 0xd7:
            body += b"\xff" * (1 + k % 3)
            k += 1
        body.append(scan[i])
        i += 1
    return data[:sc.begin] + bytes(body) + data[sc.end:], len(body) - len(scan)


def decode_and_read(data: bytes, mode: int):
    """Decode through the C ABI; returns (scan info tuple, destuffed bytes, segment directory,
    control counters, sync records)."""
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    sess = DecodeSession("cuda:0")
    sess.dec.set_deferred_walk(mode)
    host = pinned_bytes(data)
    sess.parse(host)
    planes = sess.alloc_planes()
    sess.transfer()
    sess.decode(planes)
    torch.cuda.synchronize()
    st = sess.dec.scan_stats(0)  # (runs the host walk for the exact counts; offsets do not move)
    info = sess.read_scratch(st.off_scan_info, 16).view(np.uint32).copy()
    n_subs, n_segs = int(info[0]), int(info[1])
    words = sess.read_scratch(st.off_destuffed, n_subs * 64).view(np.uint32).copy()
    destuffed = words.byteswap().view(np.uint8)  # the kernel stores big-endian words
    segs = sess.read_scratch(st.off_segments, 16 * n_segs).view(np.uint32).reshape(-1, 4).copy()
    ctrl = sess.read_scratch(st.off_control, 64).view(np.uint32).copy()
    recs = sess.read_scratch(st.off_sync_records, 16 * n_subs).view(np.uint16).reshape(-1, 8).copy()
    sess.close()
    return info, destuffed, segs, {n: int(ctrl[i]) for i, n in enumerate(CTRL_NAMES)}, recs, st


def check(data: bytes, label: str, modes=(0, 1, 3), expect_status=0, oracle_data=None, inserted=0):
    """oracle_data: the file the oracle destuffs (default: data itself); inserted: bytes `data`
    holds more than oracle_data inside the scan."""
    odata = data if oracle_data is None else oracle_data
    st, oinfo = ob.parse(odata)
    assert st == 0, (label, st)
    sc = oinfo.scans[0]
    want, seg_off, seg_len = ob.destuff(odata[sc.begin:sc.end], 64)
    for mode in modes:
        info, got, segs, ctrl, recs, stats = decode_and_read(data, mode)
        tag = f"{label} mode {mode}"
        assert int(info[0]) * 64 == len(want), tag
        assert int(info[1]) == len(seg_off), tag
        # the marker code behind the 0xff that ends the scan
        assert int(info[2]) == sc.end - sc.begin + inserted + 1, tag
        assert int(info[3]) == expect_status, tag
        assert np.array_equal(got, want), f"{tag}: {(got != want).sum()} destuffed bytes differ"
        assert np.array_equal(segs[:, 0].astype(np.int64) * 64, seg_off), tag
        assert np.array_equal(segs[:, 1].astype(np.int64), seg_len), tag
        # sync records: entry state == exit state of the predecessor, heads start from zero
        g, x = recs[:, 0], recs[:, 1]
        head = (g & 0x8000) != 0
        assert head[0] and np.all(g[head] == 0x8000), tag
        chained = ~head[1:]
        assert np.array_equal(g[1:][chained] & 0x7fff, x[:-1][chained]), tag
        assert int(head.sum()) == int((seg_len > 0).sum()), tag
        assert ctrl["slow_repairs"] == 0 and ctrl["link_rounds"] == 0, (tag, ctrl)


def test_no_restart_markers():
    check(inputs.pillow_jpeg(w=1600, h=1200, seed=31, mode="YCbCr", subsampling=2, quality=92), "no DRI")


def test_restart_markers_every_row():
    check(inputs.pillow_jpeg(w=1600, h=1200, seed=32, mode="YCbCr", subsampling=2, quality=92,
                             restart_blocks=100), "DRI 100")


def test_restart_marker_every_mcu():
    # 7 500 segments of a few bytes each: many closings per destuff tile
    check(inputs.pillow_jpeg(w=1600, h=1200, seed=33, mode="YCbCr", subsampling=2, quality=90,
                             restart_blocks=1), "DRI 1")


def test_fill_bytes_before_markers():
    plain = inputs.pillow_jpeg(w=1600, h=1200, seed=34, mode="YCbCr", subsampling=2, quality=92,
                               restart_blocks=37)
    filled, inserted = with_fill_bytes(plain)
    assert inserted > 100
    check(filled, "fill bytes", oracle_data=plain, inserted=inserted)


def test_stuffed_ff_runs():
    # quality 100 noise: long runs of 0xff 0x00 pairs, some across tile and thread boundaries
    check(inputs.pillow_jpeg(w=1200, h=900, seed=35, mode="YCbCr", subsampling=0, quality=100), "q100")


def test_bytes_behind_eoi_look_like_markers():
    data = inputs.pillow_jpeg(w=1600, h=1200, seed=36, mode="YCbCr", subsampling=2, quality=92,
                              restart_blocks=100)
    junk = (b"\xff\xd0junk\xff\x00\xff\xd9\xff\xff\xd3" + bytes(range(256))) * 400
    check(data + junk, "marker-polluted trailer")


def test_img6510_trailer_and_segments():
    data = inputs.img6510()
    if data is None:
        pytest.skip("IMG_6510.JPG not staged")
    check(data, "IMG_6510")


def test_synthetic_12mp():
    data = inputs.synth(4032, 3024, 6510)
    if data is None:
        pytest.skip("stb generator not built")
    check(data, "stb 12 MP")


def test_truncated_file_is_flagged_by_the_device():
    """No terminating marker: the host walk refuses the file at parse time (reference behaviour,
    src/reader.cpp:443-489); the device walk decodes what is there and records the fact."""
    import ctypes as C
    import jpeggpu_b200 as jb
    data = inputs.pillow_jpeg(w=1600, h=1200, seed=37, mode="YCbCr", subsampling=2, quality=92)
    cut = data[:len(data) * 3 // 4]
    if cut[-1] == 0xff:
        cut = cut[:-1]
    lib = jb.load_library()
    d = jb.Decoder(lib)
    d.set_deferred_walk(0)
    buf = np.frombuffer(cut, dtype=np.uint8)
    info = jb.ImgInfo()
    assert lib.jpeggpu_decoder_parse_header(d.handle, C.byref(info), buf.ctypes.data, len(cut)) == jb.INVALID_JPEG
    d.cleanup()
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    sess = DecodeSession("cuda:0")
    sess.parse(pinned_bytes(cut))
    planes = sess.alloc_planes()
    sess.transfer()
    sess.decode(planes)
    torch.cuda.synchronize()
    si = sess.scan_info()
    assert si.status == 1  # JPEGGPU_B200_SCAN_UNTERMINATED
    assert si.num_segments == 1 and si.num_subsequences > 0
    assert si.scan_end == len(cut) - ob.parse(data)[1].scans[0].begin  # the data ran out there
    # the rows that were there are decoded exactly as in the whole file
    ref = ob.decode(data)
    top = planes[0][: 16 * 40].cpu().numpy()[:, : ref["pixels"][0].shape[1]]
    assert np.array_equal(top, ref["pixels"][0][: 16 * 40])
    sess.close()


@pytest.mark.parametrize("which", ["stb 12 MP", "IMG_6510", "stb 48 MP", "stb 2 MP"])
def test_sync_needs_no_repair_on_any_bench_configuration(which):
    """BASELINE.json's single-image configurations: the sequential repair must not run
    (slow_repairs == 0: it is bit-exact but many times slower), and the chunks whose guess of the
    entry state had to be replaced by the left neighbour's (relinked) stay a small share: each of
    them costs its warp a second fix-point iteration."""
    data = {"stb 12 MP": lambda: inputs.synth(4032, 3024, 6510), "IMG_6510": inputs.img6510,
            "stb 48 MP": lambda: inputs.synth(8064, 6048, 48), "stb 2 MP": lambda: inputs.synth(1920, 1080, 1)}[which]()
    if data is None:
        pytest.skip("input file not staged")
    info, got, segs, ctrl, recs, st = decode_and_read(data, 3)
    chunks = (int(info[0]) + 25) // 26
    assert ctrl["slow_repairs"] == 0 and ctrl["link_rounds"] == 0, ctrl
    assert ctrl["relinked"] <= max(4, chunks // 4), (ctrl, chunks)
    g, x = recs[:, 0], recs[:, 1]
    head = (g & 0x8000) != 0
    chained = ~head[1:]
    assert np.array_equal(g[1:][chained] & 0x7fff, x[:-1][chained])
