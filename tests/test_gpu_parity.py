"""Parity of the CUDA path (through the C ABI) against the CPU oracle, bit-exact:
quantised coefficient planes and output samples.  Runs on the B200 box (-m gpu)."""
import numpy as np
import pytest

import inputs
import oracle_bindings as ob

pytestmark = pytest.mark.gpu


def run_ours(data, want_coefficients=True):
    from jpeggpu_b200.torch_io import decode_file
    return decode_file(data, want_coefficients=want_coefficients)


def check_against_oracle(data, label):
    ref = ob.decode(data)
    assert ref["status"] == 0, (label, ref["status"])
    info, planes, coefs = run_ours(data)
    oi = ref["info"]
    assert info.num_components == oi.num_components
    for c in range(info.num_components):
        assert (info.sizes_x[c], info.sizes_y[c]) == (oi.size_x[c], oi.size_y[c])
        assert coefs[c].shape == ref["planes"][c].shape, (label, c)
        ndiff = int((coefs[c] != ref["planes"][c]).sum())
        assert ndiff == 0, f"{label}: component {c}: {ndiff} coefficients differ"
        pd = int((planes[c] != ref["pixels"][c]).sum())
        assert pd == 0, f"{label}: component {c}: {pd} samples differ"


def test_img6510_matches_oracle():
    data = inputs.img6510()
    if data is None:
        pytest.skip("IMG_6510.JPG not staged (run tools/gen_inputs.py where /root/reference exists)")
    check_against_oracle(data, "IMG_6510")


@pytest.mark.parametrize("w,h,seed,q", [(4032, 3024, 6510, 90), (3008, 2000, 1, 90),
                                        (3008, 2000, 2, 95), (3001, 1999, 3, 90),
                                        (1920, 1080, 1, 90)])
def test_stb_synthetic_matches_oracle(w, h, seed, q):
    data = inputs.synth(w, h, seed, q)
    if data is None:
        pytest.skip("stb generator not built")
    check_against_oracle(data, f"stb {w}x{h} q{q}")


def test_48mp_no_restart_matches_oracle():
    data = inputs.synth(8064, 6048, 48)
    if data is None:
        pytest.skip("stb generator not built")
    check_against_oracle(data, "stb 48MP")


PILLOW_CASES = [
    dict(w=3008, h=2000, seed=11, mode="YCbCr", subsampling=0),
    dict(w=3008, h=2000, seed=12, mode="YCbCr", subsampling=1),
    dict(w=3008, h=2000, seed=13, mode="YCbCr", subsampling=2),
    dict(w=3008, h=2000, seed=14, mode="L"),
    dict(w=1504, h=1000, seed=15, mode="CMYK"),
    dict(w=1000, h=750, seed=16, mode="YCbCr", subsampling=2, restart_blocks=1),
    dict(w=1000, h=750, seed=17, mode="YCbCr", subsampling=2, restart_blocks=7),
    dict(w=1000, h=750, seed=18, mode="YCbCr", subsampling=1, restart_blocks=63),
    dict(w=1000, h=750, seed=19, mode="YCbCr", subsampling=2, optimize=True, quality=97),
    dict(w=1000, h=750, seed=20, mode="YCbCr", subsampling=0, optimize=True, quality=30),
    dict(w=8, h=8, seed=21, mode="L"),
    dict(w=1, h=1, seed=22, mode="YCbCr", subsampling=2),
    dict(w=17, h=13, seed=23, mode="YCbCr", subsampling=2),
    dict(w=250, h=3, seed=24, mode="YCbCr", subsampling=1, restart_blocks=2),
    dict(w=640, h=480, seed=25, mode="L", restart_blocks=5, quality=100),
]


@pytest.mark.parametrize("case", PILLOW_CASES, ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items()))
def test_libjpeg_encoded_variants_match_oracle(case):
    data = inputs.pillow_jpeg(**case)
    check_against_oracle(data, str(case))


def test_pitch_and_untouched_bytes():
    """Bytes between sizes_x and pitch stay untouched (reference contract)."""
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    data = inputs.pillow_jpeg(w=333, h=77, seed=31, mode="YCbCr", subsampling=2)
    ref = ob.decode(data)
    sess = DecodeSession()
    info = sess.parse(pinned_bytes(data))
    sess.transfer()
    planes = []
    for c in range(info.num_components):
        t = torch.full((info.sizes_y[c], info.sizes_x[c] + 13), 0xAB, dtype=torch.uint8, device="cuda")
        planes.append(t)
    sess.decode(planes)
    torch.cuda.synchronize()
    for c, t in enumerate(planes):
        a = t.cpu().numpy()
        assert np.array_equal(a[:, : info.sizes_x[c]], ref["pixels"][c])
        assert (a[:, info.sizes_x[c]:] == 0xAB).all()
    sess.close()


def test_decoder_reuse_and_determinism():
    """One decoder object decodes different images back to back (parse resets all state)."""
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    files = [inputs.pillow_jpeg(w=640, h=480, seed=41, mode="YCbCr", subsampling=2),
             inputs.pillow_jpeg(w=100, h=60, seed=42, mode="L"),
             inputs.pillow_jpeg(w=800, h=600, seed=43, mode="YCbCr", subsampling=0, restart_blocks=3)]
    sess = DecodeSession()
    for rep in range(2):
        for data in files:
            ref = ob.decode(data)
            info = sess.parse(pinned_bytes(data))
            sess.transfer()
            planes = sess.alloc_planes()
            sess.decode(planes)
            torch.cuda.synchronize()
            for c, t in enumerate(planes):
                assert np.array_equal(t.cpu().numpy()[:, : info.sizes_x[c]], ref["pixels"][c])
    sess.close()


@pytest.mark.parametrize("fill", [0x00, 0xFF, 0xA5, "random"])
def test_scratch_contents_do_not_matter(fill):
    """The library owns no device memory and clears what it needs itself: whatever the caller's
    scratch buffer holds before transfer (a previous image, garbage), the result is the same.
    Also repeats the decode several times on the same transfer (bench.py's device-only loop)."""
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    data = inputs.pillow_jpeg(w=1000, h=750, seed=51, mode="YCbCr", subsampling=2, restart_blocks=9)
    ref = ob.decode(data)
    sess = DecodeSession()
    host = pinned_bytes(data)
    info = sess.parse(host)
    if fill == "random":
        g = torch.Generator(device="cuda").manual_seed(7)
        sess.scratch.copy_(torch.randint(0, 256, (sess.scratch.numel(),), dtype=torch.uint8,
                                         device="cuda", generator=g))
    else:
        sess.scratch.fill_(fill)
    sess.transfer()
    planes = sess.alloc_planes()
    for rep in range(4):
        for p in planes:
            p.fill_(0x5A)
        sess.decode(planes)
        torch.cuda.synchronize()
        for c, t in enumerate(planes):
            assert np.array_equal(t.cpu().numpy()[:, : info.sizes_x[c]], ref["pixels"][c]), (fill, rep, c)
    sess.close()


def test_many_streams_concurrently():
    """Independent decoders on independent streams of one device (the batch protocol): kernels
    of different images overlap on the GPU, look-back tickets and mailboxes are per scratch."""
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    files = [inputs.pillow_jpeg(w=1000 + 16 * i, h=700, seed=60 + i, mode="YCbCr", subsampling=2,
                                restart_blocks=(0, 5, 63)[i % 3]) for i in range(6)]
    refs = [ob.decode(d) for d in files]
    streams = [torch.cuda.Stream() for _ in files]
    sessions, outs = [], []
    for data in files:
        s = DecodeSession()
        s.parse(pinned_bytes(data))
        sessions.append(s)
    torch.cuda.synchronize()
    for rep in range(3):
        outs = []
        for s, st in zip(sessions, streams):
            with torch.cuda.stream(st):
                planes = s.alloc_planes()
            st.synchronize()
            s.transfer(st)
            s.decode(planes, st)
            outs.append(planes)
        torch.cuda.synchronize()
        for s, planes, ref in zip(sessions, outs, refs):
            for c, t in enumerate(planes):
                assert np.array_equal(t.cpu().numpy()[:, : s.info.sizes_x[c]], ref["pixels"][c])
    for s in sessions:
        s.close()


@pytest.mark.parametrize("defer", [True, False], ids=["deferred_walk", "eager_walk"])
@pytest.mark.parametrize("tail", [b"", b"\xff\xd9", b"junk after eoi"], ids=["plain", "double_eoi", "trailing"])
def test_deferred_and_eager_walk_agree(defer, tail):
    """jpeggpu_b200_set_deferred_walk: the walk runs inside transfer (overlapping the upload)
    or inside parse_header; scratch filled with garbage so that nothing depends on the slack
    between the planned bounds and the exact counts.  `double_eoi`: the file ends with EOI but
    the scan ends two bytes earlier than the bound assumed."""
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    base = inputs.synth(1920, 1080, 2, 90)
    if base is None:
        pytest.skip("stb generator not built")
    data = base + tail
    ref = ob.decode(data)
    assert ref["status"] == 0
    sess = DecodeSession()
    sess.dec.set_deferred_walk(defer)
    for rep in range(2):
        info = sess.parse(pinned_bytes(data))
        sess.scratch.fill_(0xC3)
        sess.transfer()
        planes = sess.alloc_planes()
        sess.decode(planes)
        coefs = sess.decode_coefficients()
        torch.cuda.synchronize()
        for c, t in enumerate(planes):
            assert np.array_equal(t.cpu().numpy()[:, : info.sizes_x[c]], ref["pixels"][c]), (rep, c)
            assert np.array_equal(coefs[c].cpu().numpy(), ref["planes"][c]), (rep, c)
    sess.close()


def _run_with_variant(name, define, code):
    """Run `code` in a subprocess against a variant of the library (jpeggpu_b200/csrc/Makefile:
    variants; built here with nvcc if build() has not)."""
    import os
    import shutil
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    csrc = os.path.join(root, "jpeggpu_b200", "csrc")
    out = os.path.join(root, "jpeggpu_b200", "lib", "variants", name)
    if not os.path.exists(out):
        if shutil.which("nvcc") is None:
            pytest.skip("variant library not built and nvcc not available")
        os.makedirs(os.path.dirname(out), exist_ok=True)
        srcs = ["kernels_destuff.cu", "kernels_huffman.cu", "kernels_dc.cu", "kernels_idct.cu",
                "decoder.cpp", "parser.cpp", "batch.cpp"]
        subprocess.run(["nvcc", "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a",
                        "-Xcompiler", "-fPIC,-O3", "-I" + os.path.join(root, "include"), "-I" + csrc,
                        define, "-shared", "-Xlinker", "-Bsymbolic", "-x", "cu",
                        *[os.path.join(csrc, f) for f in srcs], "-lpthread", "-o", out], check=True)
    env = dict(os.environ, JPEGGPU_B200_LIB=out,
               PYTHONPATH=os.pathsep.join([root, os.path.join(root, "tests"), os.path.join(root, "tools")]))
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-2000:]
    return r.stdout


VARIANT_DECODES = (
    "import numpy as np, inputs, oracle_bindings as ob\n"
    "from jpeggpu_b200.torch_io import decode_file\n"
    "for kw in (dict(w=1600, h=1200, seed=81, mode='YCbCr', subsampling=2, quality=92),\n"
    "           dict(w=1200, h=900, seed=82, mode='YCbCr', subsampling=0, quality=100),\n"
    "           dict(w=1280, h=720, seed=83, mode='YCbCr', subsampling=2, restart_blocks=40)):\n"
    "    data = inputs.pillow_jpeg(**kw)\n"
    "    ref = ob.decode(data)\n"
    "    info, planes, coefs = decode_file(data, want_coefficients=True)\n"
    "    for c in range(info.num_components):\n"
    "        assert np.array_equal(coefs[c], ref['planes'][c]), (kw, c)\n"
    "        assert np.array_equal(planes[c], ref['pixels'][c]), (kw, c)\n")


def test_write_list_overflow_path_is_bit_exact():
    """huff_write keeps a bounded list of coefficients per thread (32 entries) and sends the
    overflow straight to global memory.  Both paths run on ordinary images; this test shifts the
    balance with a library built with 8-entry lists (-DJB2_WRITE_LIST=8) in a subprocess, so that
    nearly every coefficient takes the overflow path."""
    _run_with_variant("lib_list8.so", "-DJB2_WRITE_LIST=8", VARIANT_DECODES + "print('ok')\n")


def test_verify_and_repair_path_is_bit_exact():
    """huff_sync's last CTA verifies (and repairs) the chunk boundaries only when a chunk asks for
    it: its exit state changed after the right neighbour may have read it, or the wait for the
    left neighbour timed out.  On ordinary runs nobody asks (need_verify == 0, asserted in
    test_gpu_destuff.py).  A library whose mailbox wait gives up at once (-DJB2_MAILBOX_SPINS=1)
    makes most chunks go on with their own guess: the verification then has real work, and the
    repaired records must give the oracle's coefficients.  The control block must show that the
    path ran."""
    code = VARIANT_DECODES + (
        "from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes\n"
        "import torch\n"
        "data = inputs.pillow_jpeg(w=1600, h=1200, seed=84, mode='YCbCr', subsampling=2, quality=95)\n"
        "sess = DecodeSession('cuda:0'); sess.parse(pinned_bytes(data)); planes = sess.alloc_planes()\n"
        "sess.transfer(); sess.decode(planes); torch.cuda.synchronize()\n"
        "st = sess.dec.scan_stats(0)\n"
        "ctrl = sess.read_scratch(st.off_control, 64).view(np.uint32)\n"
        "ref = ob.decode(data)\n"
        "for c, t in enumerate(planes):\n"
        "    assert np.array_equal(t.cpu().numpy()[:, :ref['pixels'][c].shape[1]], ref['pixels'][c]), c\n"
        "print('ctrl', int(ctrl[0]), int(ctrl[1]), int(ctrl[2]), int(ctrl[4]))\n"
        "assert int(ctrl[4]) == 1, 'no chunk asked for the verification'\n"
        "print('ok')\n")
    out = _run_with_variant("lib_spin1.so", "-DJB2_MAILBOX_SPINS=1", code)
    print(out)


def _poisoned_decode(data, walk_mode=3):
    """Decode with every byte of the scratch buffer set to garbage first (the transfer then writes
    what it uploads): whatever the kernels read must have been written by this decode."""
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    sess = DecodeSession()
    sess.dec.set_deferred_walk(walk_mode)
    info = sess.parse(pinned_bytes(data))
    sess.scratch.fill_(0xC3)
    sess.transfer()
    planes = sess.alloc_planes()
    sess.decode(planes)
    coefs = sess.decode_coefficients()
    torch.cuda.synchronize()
    out = ([t.cpu().numpy()[:, : info.sizes_x[c]] for c, t in enumerate(planes)], [c.cpu().numpy() for c in coefs])
    sess.close()
    return out


POISON_CASES = {
    "no_dri": dict(w=1600, h=1200, seed=101, mode="YCbCr", subsampling=2, quality=92),
    "dri_1": dict(w=800, h=600, seed=102, mode="YCbCr", subsampling=2, quality=90, restart_blocks=1),
    "dri_37": dict(w=1600, h=1200, seed=103, mode="YCbCr", subsampling=1, quality=95, restart_blocks=37),
    "grey_tiny": dict(w=40, h=24, seed=104, mode="L"),
    "444_q100": dict(w=1200, h=900, seed=105, mode="YCbCr", subsampling=0, quality=100),
    "flat_q20": dict(w=2048, h=1536, seed=106, mode="YCbCr", subsampling=2, quality=20),
}


@pytest.mark.parametrize("name", list(POISON_CASES))
@pytest.mark.parametrize("walk_mode", [3, 0], ids=["device_walk", "host_walk"])
def test_nothing_is_read_that_this_decode_did_not_write(name, walk_mode):
    """huff_write clears exactly the stretch of the coefficient buffer each warp scatters into
    (plus what nobody writes); a stretch left out would show up as 0xC3C3 coefficients."""
    data = inputs.pillow_jpeg(**POISON_CASES[name])
    ref = ob.decode(data)
    assert ref["status"] == 0
    planes, coefs = _poisoned_decode(data, walk_mode)
    for c in range(len(planes)):
        assert np.array_equal(coefs[c], ref["planes"][c]), (name, c)
        assert np.array_equal(planes[c], ref["pixels"][c]), (name, c)


def test_restart_segments_without_data_are_cleared():
    """Two restart markers back to back leave a segment without a single subsequence: its data
    units hold zero coefficients (samples 128) whatever the scratch held before, like in the
    reference, which clears the whole buffer first.  (The oracle reads an empty segment as a run
    of one bits, so the expectation is built from the intact file: 640x480 4:2:0 with one MCU row
    per restart segment - the other rows decode as before, DC prediction restarts per segment.)"""
    data = inputs.pillow_jpeg(w=640, h=480, seed=107, mode="YCbCr", subsampling=2, quality=90, restart_blocks=40)
    st, info = ob.parse(data)
    sc = info.scans[0]
    scan = bytearray(data[sc.begin:sc.end])
    # drop the entropy-coded bytes of the third and fourth restart segment (keep their markers)
    marks = [i for i in range(len(scan) - 1) if scan[i] == 0xff and 0xd0 <= scan[i + 1] <= 0xd7]
    assert len(marks) == 29
    cut = bytes(scan[:marks[1] + 2]) + bytes(scan[marks[2]:marks[2] + 2]) + bytes(scan[marks[3]:])
    damaged = data[:sc.begin] + cut + data[sc.end:]
    ref = ob.decode(data)
    assert ref["status"] == 0
    for mode in (3, 0):
        planes, coefs = _poisoned_decode(damaged, mode)
        for c in range(len(planes)):
            rows = slice(32, 64) if c == 0 else slice(16, 32)  # MCU rows 2 and 3
            want_c = ref["planes"][c].copy()
            want_c[rows] = 0
            want_p = ref["pixels"][c].copy()
            want_p[rows] = 128
            assert np.array_equal(coefs[c], want_c), (mode, c)
            assert np.array_equal(planes[c], want_p), (mode, c)


def test_truncated_scan_leaves_zero_coefficients_behind():
    """A file cut in the middle of the scan: the data units behind the cut are zero (level 128 /
    the last DC), not what the previous decode left in the scratch."""
    data = inputs.pillow_jpeg(w=1600, h=1200, seed=108, mode="YCbCr", subsampling=2, quality=92)
    cut = data[: len(data) * 2 // 3]
    if cut[-1] == 0xff:
        cut = cut[:-1]
    a_planes, a_coefs = _poisoned_decode(cut, 3)
    # the same decode over a scratch poisoned with another pattern gives the same output
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    sess = DecodeSession()
    info = sess.parse(pinned_bytes(cut))
    sess.scratch.fill_(0x5A)
    sess.transfer()
    planes = sess.alloc_planes()
    sess.decode(planes)
    coefs = sess.decode_coefficients()
    torch.cuda.synchronize()
    for c in range(info.num_components):
        assert np.array_equal(coefs[c].cpu().numpy(), a_coefs[c]), c
        assert np.array_equal(planes[c].cpu().numpy()[:, : info.sizes_x[c]], a_planes[c]), c
    # and the tail of the luma coefficient plane holds no AC coefficient (the DC of every data
    # unit is the last predictor: nothing restarts it)
    tail = a_coefs[0][-64:].copy()
    tail[::8, ::8] = 0
    assert not tail.any()
    sess.close()


def test_writers_clear_their_own_stretches_variant():
    """Above 100 MB of coefficients (48 MP) huff_write clears the buffer itself, stretch by stretch;
    below, huff_sync clears it up front.  A library built with the threshold at zero
    (-DJB2_FILL_IN_WRITE_BYTES=0) runs the first path on the small poisoned inputs of this file."""
    code = (
        "import numpy as np, inputs, oracle_bindings as ob\n"
        "import test_gpu_parity as t\n"
        "for name, kw in t.POISON_CASES.items():\n"
        "    data = inputs.pillow_jpeg(**kw)\n"
        "    ref = ob.decode(data)\n"
        "    for mode in (3, 0):\n"
        "        planes, coefs = t._poisoned_decode(data, mode)\n"
        "        for c in range(len(planes)):\n"
        "            assert np.array_equal(coefs[c], ref['planes'][c]), (name, mode, c)\n"
        "            assert np.array_equal(planes[c], ref['pixels'][c]), (name, mode, c)\n"
        "t.test_restart_segments_without_data_are_cleared()\n"
        "t.test_truncated_scan_leaves_zero_coefficients_behind()\n"
        "print('ok')\n")
    _run_with_variant("lib_fillwrite.so", "-DJB2_FILL_IN_WRITE_BYTES=0", code)
