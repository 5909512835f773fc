"""GPU parity of the optional RGB output through the C ABI, bit-exact against
the oracle's restatement of the reference's conv_to_rgbi (util/util.h:33-107):
jpeggpu_b200_decode_rgbi (fused into the IDCT kernel) and jpeggpu_b200_convert_rgbi (planes that
already exist).  The oracle itself is pinned against the reference function in
tests/test_rgbi_oracle.py."""
import numpy as np
import pytest

import inputs
import oracle_bindings as ob
import test_rgbi_oracle as cases

pytestmark = pytest.mark.gpu


def oracle_rgb(data):
    ref = ob.decode(data)
    assert ref["status"] == 0
    oi = ref["info"]
    n = oi.num_components
    rc, rgb = ob.conv_to_rgbi(ref["pixels"], list(oi.ssx)[:n], list(oi.ssy)[:n])
    assert rc == 0
    return rgb


def fused_rgb(data, pitch_extra=0, fill=0x5a):
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    sess = DecodeSession("cuda:0")
    info = sess.parse(pinned_bytes(data))
    sess.transfer()
    h, w = info.sizes_y[0], info.sizes_x[0]
    buf = torch.full((h, 3 * w + pitch_extra), fill, dtype=torch.uint8, device="cuda:0")
    sess.dec.decode_rgbi(buf.data_ptr(), buf.stride(0), sess.scratch.data_ptr(), sess.scratch.numel(),
                         torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    out = buf.cpu().numpy()
    sess.close()
    return out[:, : 3 * w].reshape(h, w, 3), out[:, 3 * w:]


RGB_CASES = [
    dict(w=1280, h=720, seed=31, mode="YCbCr", subsampling=2),
    dict(w=1000, h=750, seed=32, mode="YCbCr", subsampling=1),
    dict(w=1000, h=750, seed=33, mode="YCbCr", subsampling=0),
    dict(w=1001, h=747, seed=34, mode="YCbCr", subsampling=2, restart_blocks=7),
    dict(w=999, h=333, seed=35, mode="YCbCr", subsampling=1, optimize=True, quality=97),
    dict(w=515, h=389, seed=36, mode="YCbCr", subsampling=0, quality=30),
    dict(w=3008, h=2000, seed=37, mode="L"),
    dict(w=643, h=481, seed=38, mode="L", restart_blocks=5, quality=100),
    dict(w=1, h=1, seed=39, mode="YCbCr", subsampling=2),
    dict(w=17, h=13, seed=40, mode="YCbCr", subsampling=2),
    dict(w=250, h=3, seed=41, mode="YCbCr", subsampling=1, restart_blocks=2),
    dict(w=8, h=8, seed=42, mode="L"),
]


@pytest.mark.parametrize("case", RGB_CASES, ids=lambda c: "-".join(f"{k}{v}" for k, v in c.items()))
def test_decode_rgbi_matches_oracle(case):
    data = inputs.pillow_jpeg(**case)
    want = oracle_rgb(data)
    got, _ = fused_rgb(data)
    assert got.shape == want.shape
    nd = int((got != want).sum())
    assert nd == 0, f"{case}: {nd} RGB bytes differ"


@pytest.mark.parametrize("extra", [5, 8, 64])
def test_decode_rgbi_pitch_and_untouched_bytes(extra):
    """Odd pitches take the byte-store path; bytes beyond 3*width of a row are never written."""
    data = inputs.pillow_jpeg(w=333, h=222, seed=43, mode="YCbCr", subsampling=2)
    want = oracle_rgb(data)
    got, pad = fused_rgb(data, pitch_extra=extra, fill=0xa7)
    assert np.array_equal(got, want)
    assert (pad == 0xa7).all()


def test_decode_rgbi_12mp_workload_and_photo():
    for data in (inputs.synth(4032, 3024, 6510), inputs.img6510()):
        if data is None:
            pytest.skip("input not staged")
        want = oracle_rgb(data)
        got, _ = fused_rgb(data)
        assert np.array_equal(got, want)


def test_decode_rgbi_then_planes_on_one_transfer():
    """Both outputs can be produced from one transfer; planes are unaffected by the RGB pass."""
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    data = inputs.pillow_jpeg(w=640, h=480, seed=44, mode="YCbCr", subsampling=2)
    ref = ob.decode(data)
    sess = DecodeSession("cuda:0")
    sess.parse(pinned_bytes(data))
    sess.transfer()
    rgb = sess.decode_rgbi()
    planes = sess.alloc_planes()
    sess.decode(planes)
    rgb2 = sess.decode_rgbi()
    torch.cuda.synchronize()
    assert np.array_equal(rgb.cpu().numpy(), oracle_rgb(data))
    assert torch.equal(rgb, rgb2)
    for c in range(3):
        assert np.array_equal(planes[c].cpu().numpy(), ref["pixels"][c])
    sess.close()


def _convert(planes, fx, fy, pitch_extra=0):
    import torch
    import jpeggpu_b200 as jb
    info = jb.ImgInfo()
    info.num_components = len(planes)
    dev = []
    for c, p in enumerate(planes):
        info.sizes_x[c], info.sizes_y[c] = p.shape[1], p.shape[0]
        info.subsampling.x[c] = fx if c == 0 else 1
        info.subsampling.y[c] = fy if c == 0 else 1
        t = torch.zeros((p.shape[0], p.shape[1] + (3 if pitch_extra else 0)), dtype=torch.uint8, device="cuda:0")
        t[:, : p.shape[1]] = torch.from_numpy(np.ascontiguousarray(p)).cuda()
        dev.append(t)
    h, w = planes[0].shape
    out = torch.zeros((h, 3 * w + pitch_extra), dtype=torch.uint8, device="cuda:0")
    jb.convert_rgbi([t.data_ptr() for t in dev], [t.stride(0) for t in dev], info, out.data_ptr(),
                    out.stride(0), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return out.cpu().numpy()[:, : 3 * w].reshape(h, w, 3)


def test_convert_rgbi_every_triple():
    """All 2^24 (y, cb, cr) inputs: the device arithmetic equals the oracle's everywhere."""
    planes = cases.all_triples()
    rc, want = ob.conv_to_rgbi(planes, [1, 1, 1], [1, 1, 1])
    assert rc == 0
    got = _convert(planes, 1, 1)
    bad = np.argwhere((got != want).any(axis=2))
    assert bad.size == 0, f"{len(bad)} triples differ, first (y,cb,cr) index {bad[:3].tolist()}"


@pytest.mark.parametrize("w,h,fx,fy", cases.CASES)
@pytest.mark.parametrize("extra", [0, 5])
def test_convert_rgbi_subsampled(w, h, fx, fy, extra):
    planes = cases.subsampled_case(w, h, fx, fy, seed=w * 131 + h)
    rc, want = ob.conv_to_rgbi(planes, [fx, 1, 1], [fy, 1, 1])
    assert rc == 0
    assert np.array_equal(_convert(planes, fx, fy, pitch_extra=extra), want)


def test_convert_rgbi_grey_and_multiscan_route():
    """A three-scan file is refused by decode_rgbi; decode + convert_rgbi gives the same picture
    as the oracle."""
    import torch
    import jpeg_transcode as jt
    import jpeggpu_b200 as jb
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    src = inputs.pillow_jpeg(w=320, h=200, seed=45, mode="YCbCr", subsampling=2)
    multi = jt.transcode(src, ob.decode(src)["planes"], scans=[[0], [1], [2]])
    want = oracle_rgb(multi)
    sess = DecodeSession("cuda:0")
    info = sess.parse(pinned_bytes(multi))
    sess.transfer()
    with pytest.raises(jb.JpegGpuError) as e:
        sess.decode_rgbi()
    assert e.value.status == jb.NOT_SUPPORTED
    planes = sess.alloc_planes()
    sess.decode(planes)
    out = torch.zeros((200, 320, 3), dtype=torch.uint8, device="cuda:0")
    jb.convert_rgbi([p.data_ptr() for p in planes], [p.stride(0) for p in planes], info, out.data_ptr(),
                    out.stride(0), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), want)
    sess.close()
    g = cases.subsampled_case(77, 31, 1, 1, 9)[:1]
    assert np.array_equal(_convert(g, 1, 1), np.repeat(g[0][:, :, None], 3, axis=2))
