"""Fused batches (include/jpeggpu_b200.h: jpeggpu_b200_batch_*): K images per kernel launch.
The reference has no batch entry point (its callers loop over the per-image protocol,
benchmark/benchmark_jpeggpu.hpp:69-89); the outputs must be what that loop produces, i.e. what the
oracle produces per file.  Host-only planning is covered without a GPU; the decode is -m gpu."""
import ctypes as C

import numpy as np
import pytest

import inputs
import oracle_bindings as ob


def _lib():
    import jpeggpu_b200 as jb
    return jb, jb.load_library()


def _mixed_files():
    return [
        inputs.pillow_jpeg(w=640, h=480, seed=41, mode="YCbCr", subsampling=2),
        inputs.pillow_jpeg(w=803, h=601, seed=42, mode="YCbCr", subsampling=0, quality=95),
        inputs.pillow_jpeg(w=1280, h=720, seed=43, mode="YCbCr", subsampling=2, restart_blocks=40),
        inputs.pillow_jpeg(w=512, h=512, seed=44, mode="L"),
        inputs.pillow_jpeg(w=333, h=211, seed=45, mode="YCbCr", subsampling=1, optimize=True),
        inputs.pillow_jpeg(w=64, h=48, seed=46, mode="YCbCr", subsampling=2),
        inputs.pillow_jpeg(w=1600, h=1200, seed=47, mode="YCbCr", subsampling=2, quality=92),
        inputs.pillow_jpeg(w=500, h=375, seed=48, mode="YCbCr", subsampling=2, restart_blocks=1),
    ]


def test_plan_is_host_only_and_reports_per_item_status():
    jb, lib = _lib()
    from jpeggpu_b200.batch import FusedBatch, make_items
    files = _mixed_files()[:4]
    multi = inputs.pillow_jpeg(w=320, h=240, seed=5, mode="CMYK")  # Pillow: one scan of four components
    bad = files[0][:100]
    bufs = [np.frombuffer(f, dtype=np.uint8) for f in files + [bad]]
    items = make_items([(b.ctypes.data, b.size) for b in bufs] + [(None, 0)])
    fb = FusedBatch(lib)
    need, st = fb.plan(items, strict=False)
    assert need > 0
    assert st != jb.SUCCESS
    for i, f in enumerate(files):
        ost, oi = ob.parse(f)
        assert items[i].status == jb.SUCCESS
        info = items[i].info
        assert info.num_components == oi.num_components
        for c in range(info.num_components):
            assert (info.sizes_x[c], info.sizes_y[c]) == (oi.size_x[c], oi.size_y[c])
    assert items[len(files)].status != jb.SUCCESS          # truncated inside the headers
    assert items[len(files) + 1].status == jb.INVALID_ARGUMENT
    # planning the same batch again gives the same arena; a smaller batch a smaller one
    need2, _ = fb.plan(items, strict=False)
    assert need2 == need
    need3, st3 = fb.plan(items, count=2)
    assert st3 == jb.SUCCESS and 0 < need3 < need
    fb.close()
    del multi


def _decode_fused(files, contiguous, repeats=1):
    """Returns per file the list of output planes (numpy) of the last repeat."""
    import torch
    jb, lib = _lib()
    from jpeggpu_b200.batch import FusedBatch, make_items
    if contiguous:
        # the caller's "pinned arena": all files back to back, at odd offsets on purpose
        total = sum(len(f) + 7 for f in files)
        host = torch.empty(total, dtype=torch.uint8, pin_memory=True)
        view = host.numpy()
        ptrs, at = [], 3
        for f in files:
            view[at:at + len(f)] = np.frombuffer(f, dtype=np.uint8)
            ptrs.append((host.data_ptr() + at, len(f)))
            at += len(f) + 7
        keep = [host]
        arena_range = (host.data_ptr(), host.numel())
    else:
        arena_range = None
        keep, ptrs = [], []
        for f in files:
            t = torch.empty(len(f) + 64 * 1024 * 1024 // 8, dtype=torch.uint8, pin_memory=True)  # far apart
            t.numpy()[:len(f)] = np.frombuffer(f, dtype=np.uint8)
            keep.append(t)
            ptrs.append((t.data_ptr(), len(f)))
    items = make_items(ptrs)
    fb = FusedBatch(lib)
    if arena_range:
        fb.set_host_arena(*arena_range)
    need, st = fb.plan(items)
    arena = torch.empty(need, dtype=torch.uint8, device="cuda:0")
    planes = []
    for it in items:
        ps = [torch.full((it.info.sizes_y[c], it.info.sizes_x[c] + 5), 0xAB, dtype=torch.uint8, device="cuda:0")
              for c in range(it.info.num_components)]
        for c, p in enumerate(ps):
            it.img.image[c] = p.data_ptr()
            it.img.pitch[c] = p.stride(0)
        planes.append(ps)
    stream = torch.cuda.current_stream().cuda_stream
    for r in range(repeats):
        if r:
            need_r, _ = fb.plan(items)
            assert need_r == need
        fb.decode(items, arena.data_ptr(), arena.numel(), stream)
    torch.cuda.synchronize()
    infos = [fb.read_scan_info(i, arena.data_ptr(), stream) for i in range(len(files))]
    out = [[p.cpu().numpy() for p in ps] for ps in planes]
    fb.close()
    del keep
    return items, out, infos


@pytest.mark.gpu
@pytest.mark.parametrize("contiguous", [True, False], ids=["one_span_copy", "copy_per_file"])
def test_mixed_geometry_batch_is_bit_exact(contiguous):
    files = _mixed_files()
    items, out, infos = _decode_fused(files, contiguous)
    for i, f in enumerate(files):
        ref = ob.decode(f)
        assert ref["status"] == 0
        st, oinfo = ob.parse(f)
        sc = oinfo.scans[0]
        assert infos[i].status == 0 and infos[i].scan_end == sc.end - sc.begin + 1, i
        for c, want in enumerate(ref["pixels"]):
            got = out[i][c]
            assert np.array_equal(got[:, :want.shape[1]], want), f"file {i} component {c}"
            assert np.all(got[:, want.shape[1]:] == 0xAB), "bytes between sizes_x and pitch were touched"


@pytest.mark.gpu
def test_batch_object_is_reusable_and_matches_the_per_image_path():
    import torch
    from jpeggpu_b200.torch_io import decode_file
    files = [inputs.pillow_jpeg(w=1920, h=1080, seed=60 + k, mode="YCbCr", subsampling=2, quality=90)
             for k in range(12)]
    items, out, infos = _decode_fused(files, True, repeats=3)
    for i, f in enumerate(files):
        info, planes, _ = decode_file(f)
        for c in range(info.num_components):
            assert np.array_equal(out[i][c][:, :planes[c].shape[1]], planes[c]), (i, c)


@pytest.mark.gpu
def test_items_the_fused_path_cannot_take_are_skipped():
    """A damaged file inside a batch: its status says so, the other items decode."""
    import torch
    jb, lib = _lib()
    from jpeggpu_b200.batch import FusedBatch, make_items
    good = inputs.pillow_jpeg(w=640, h=480, seed=71, mode="YCbCr", subsampling=2)
    files = [good, good[:200], good]
    hosts = []
    for f in files:
        t = torch.empty(len(f), dtype=torch.uint8, pin_memory=True)
        t.numpy()[:] = np.frombuffer(f, dtype=np.uint8)
        hosts.append(t)
    items = make_items([(t.data_ptr(), t.numel()) for t in hosts])
    fb = FusedBatch(lib)
    need, st = fb.plan(items, strict=False)
    assert st != jb.SUCCESS
    assert items[0].status == jb.SUCCESS and items[2].status == jb.SUCCESS and items[1].status != jb.SUCCESS
    arena = torch.empty(need, dtype=torch.uint8, device="cuda:0")
    planes = {}
    for i in (0, 2):
        it = items[i]
        ps = [torch.zeros((it.info.sizes_y[c], it.info.sizes_x[c]), dtype=torch.uint8, device="cuda:0")
              for c in range(it.info.num_components)]
        for c, p in enumerate(ps):
            it.img.image[c] = p.data_ptr()
            it.img.pitch[c] = p.stride(0)
        planes[i] = ps
    fb.decode(items, arena.data_ptr(), arena.numel(), torch.cuda.current_stream().cuda_stream, strict=False)
    torch.cuda.synchronize()
    ref = ob.decode(good)
    for i in (0, 2):
        for c, want in enumerate(ref["pixels"]):
            assert np.array_equal(planes[i][c].cpu().numpy(), want)
    fb.close()
