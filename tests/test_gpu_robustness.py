"""Malformed input must be contained (production use: files arrive from the network).  For
corrupted, truncated or marker-polluted entropy-coded data the library may return an error from
parse_header / transfer, or decode garbage samples, but it must return, must not write outside
the caller's planes or beyond the scratch size it asked for, must not leave a sticky CUDA error,
and the same decoder must decode the next good file bit-exactly.  (The reference makes no such
promise: SURVEY 9.1 Q5 describes out-of-bounds coefficient stores behind its zero padding.)"""
import numpy as np
import pytest

import inputs
import oracle_bindings as ob

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(180)]

GUARD = 4096
FILL = 0xCD


def scan_range(data):
    st, info = ob.parse(data)
    assert st == 0
    return info.scans[0].begin, info.scans[info.num_scans - 1].end


def mutate(data, kind, seed):
    rng = np.random.RandomState(seed)
    b = bytearray(data)
    lo, hi = scan_range(data)
    n = hi - lo
    if kind == "flips":
        for i in rng.randint(lo, hi, size=max(4, n // 500)):
            b[i] ^= 1 << rng.randint(8)
    elif kind == "random_bytes":
        for i in rng.randint(lo, hi, size=max(8, n // 100)):
            b[i] = rng.randint(256)
    elif kind == "markers":
        for i in rng.randint(lo, hi - 1, size=24):
            b[i] = 0xff
            b[i + 1] = [0xd0, 0xd3, 0xd7, 0x00, 0xff, 0x01][rng.randint(6)]
    elif kind == "zeros":
        a = lo + n // 3
        b[a: a + n // 3] = bytes(n // 3)
    elif kind == "ones":
        a = lo + n // 4
        for i in range(a, a + n // 4, 2):  # 0xff 0x00 pairs: legal stuffing, all-one bits
            b[i], b[i + 1] = 0xff, 0x00
    elif kind == "truncated":
        b = b[: lo + n // 2]
    elif kind == "truncated_eoi":
        b = b[: lo + n // 2] + b"\xff\xd9"
    elif kind == "early_eoi":
        b[lo + n // 2: lo + n // 2 + 2] = b"\xff\xd9"
    elif kind == "shuffled":
        blk = 512
        chunks = [bytes(b[i: i + blk]) for i in range(lo, hi - blk, blk)]
        rng.shuffle(chunks)
        b[lo: lo + blk * len(chunks)] = b"".join(chunks)
    return bytes(b)


KINDS = ["flips", "random_bytes", "markers", "zeros", "ones", "truncated", "truncated_eoi", "early_eoi",
         "shuffled"]
SOURCES = {
    "420": dict(w=640, h=480, seed=51, mode="YCbCr", subsampling=2),
    "420_dri": dict(w=640, h=480, seed=52, mode="YCbCr", subsampling=2, restart_blocks=9),
    "444_opt": dict(w=333, h=222, seed=53, mode="YCbCr", subsampling=0, optimize=True),
    "grey": dict(w=500, h=300, seed=54, mode="L"),
}


def guarded(shape_bytes, device):
    import torch
    buf = torch.full((GUARD + shape_bytes + GUARD,), FILL, dtype=torch.uint8, device=device)
    return buf, buf[GUARD: GUARD + shape_bytes]


def guards_intact(buf, n):
    return bool((buf[:GUARD] == FILL).all().item()) and bool((buf[GUARD + n:] == FILL).all().item())


@pytest.mark.parametrize("src", list(SOURCES))
@pytest.mark.parametrize("kind", KINDS)
def test_malformed_scan_is_contained(src, kind):
    import torch
    import jpeggpu_b200 as jb
    from jpeggpu_b200.torch_io import pinned_bytes
    good = inputs.pillow_jpeg(**SOURCES[src])
    want = ob.decode(good)
    dev = torch.device("cuda:0")
    dec = jb.Decoder()
    stream = torch.cuda.current_stream().cuda_stream
    for seed in range(3):
        bad = mutate(good, kind, seed)
        host = pinned_bytes(bad)
        try:
            info = dec.parse_header(host.data_ptr(), host.numel(), keepalive=host)
        except jb.JpegGpuError:
            continue  # refused on the host: fine
        need = dec.get_buffer_size()
        sbuf, scratch = guarded((need + 255) // 256 * 256, dev)
        # the scratch view must be 256-byte aligned like a cudaMalloc'ed buffer
        off = (-scratch.data_ptr()) % 256
        scratch = scratch[off: off + need]
        planes, bufs = [], []
        for c in range(info.num_components):
            b, p = guarded(info.sizes_x[c] * info.sizes_y[c], dev)
            bufs.append((b, info.sizes_x[c] * info.sizes_y[c]))
            planes.append(p)
        try:
            dec.transfer(scratch.data_ptr(), need, stream)
            dec.decode([p.data_ptr() for p in planes], [info.sizes_x[c] for c in range(info.num_components)],
                       scratch.data_ptr(), need, stream)
            if info.num_components in (1, 3):
                rb, rgb = guarded(3 * info.sizes_x[0] * info.sizes_y[0], dev)
                try:
                    dec.decode_rgbi(rgb.data_ptr(), 3 * info.sizes_x[0], scratch.data_ptr(), need, stream)
                except jb.JpegGpuError as e:
                    assert e.status == jb.NOT_SUPPORTED
                torch.cuda.synchronize()
                assert guards_intact(rb, 3 * info.sizes_x[0] * info.sizes_y[0]), (kind, seed, "rgb guard")
        except jb.JpegGpuError as e:
            assert e.status in (jb.INVALID_JPEG, jb.NOT_SUPPORTED, jb.INCOMPLETE_BITSTREAM), (kind, seed, e)
        torch.cuda.synchronize()  # raises on a sticky error (illegal address, launch failure)
        for b, n in bufs:
            assert guards_intact(b, n), (kind, seed, "plane guard")
        lead = sbuf[: GUARD + off]
        tail = sbuf[GUARD + off + need:]
        assert bool((lead == FILL).all().item()) and bool((tail == FILL).all().item()), (kind, seed, "scratch guard")
    # the same decoder, next file: bit-exact
    host = pinned_bytes(good)
    info = dec.parse_header(host.data_ptr(), host.numel(), keepalive=host)
    need = dec.get_buffer_size()
    scratch = torch.empty(need, dtype=torch.uint8, device=dev)
    planes = [torch.zeros((info.sizes_y[c], info.sizes_x[c]), dtype=torch.uint8, device=dev)
              for c in range(info.num_components)]
    dec.transfer(scratch.data_ptr(), need, stream)
    dec.decode([p.data_ptr() for p in planes], [p.stride(0) for p in planes], scratch.data_ptr(), need, stream)
    torch.cuda.synchronize()
    for c in range(info.num_components):
        assert np.array_equal(planes[c].cpu().numpy(), want["pixels"][c]), (kind, c)
    dec.cleanup()
