"""Multi-GPU row: a batch is partitioned across ranks round-robin, every rank decodes
its share on its own GPU, no collective on the data path.  CPU part: the partition, the native
multi-threaded batch parse and the world_size-2 bookkeeping gather on gloo.  GPU part: the native
batch decode against the oracle."""
import os
import socket

import numpy as np
import pytest

import inputs
import oracle_bindings as ob
from jpeggpu_b200 import batch as jb

NAMES = ["pil_420_rst1_333x211", "pil_420_opt_1280x720", "pil_444_1504x1000",
         "pil_420_rst63_1008x752", "stb_1920x1080_s1_q90"]


def _files():
    out = []
    for n in NAMES:
        d = inputs.read_named(n)
        if d is not None:
            out.append((n, d))
    out.append(("gray", inputs.pillow_jpeg(w=640, h=480, seed=4, mode="L")))
    out.append(("cmyk", inputs.pillow_jpeg(w=320, h=240, seed=5, mode="CMYK")))
    return out


@pytest.mark.parametrize("n,world", [(0, 1), (1, 4), (7, 2), (4096, 8), (13, 5)])
def test_partition_is_a_disjoint_cover(n, world):
    seen = []
    for r in range(world):
        part = jb.partition(n, r, world)
        assert all(i % world == r for i in part)
        seen += part
    assert sorted(seen) == list(range(n))
    sizes = [len(jb.partition(n, r, world)) for r in range(world)]
    assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        jb.partition(n, world, world)


def _parse_share(files, rank, world):
    mine = jb.partition(len(files), rank, world)
    bufs = [np.frombuffer(files[i][1], dtype=np.uint8) for i in mine]
    items = jb.make_items([(b.ctypes.data, b.size) for b in bufs])
    need = jb.parse_batch(items, threads=3) if len(mine) else 0
    return {"rank": rank, "items": len(mine), "need": need,
            "sizes": {files[i][0]: (items[k].info.sizes_x[0], items[k].info.sizes_y[0],
                                    items[k].info.num_components) for k, i in enumerate(mine)}}


def test_parse_batch_matches_oracle_geometry():
    files = _files()
    got = _parse_share(files, 0, 1)
    assert got["items"] == len(files) and got["need"] > 0
    for name, data in files:
        st, info = ob.parse(data)
        assert st == 0
        assert got["sizes"][name] == (info.size_x[0], info.size_y[0], info.num_components)


def test_parse_batch_reports_bad_items():
    import ctypes as C
    from jpeggpu_b200 import JpegGpuError, load_library
    good = np.frombuffer(inputs.pillow_jpeg(w=64, h=64, seed=1), dtype=np.uint8)
    bad = np.frombuffer(b"\xff\xd8\xff\xdb\x00", dtype=np.uint8)
    items = jb.make_items([(good.ctypes.data, good.size), (bad.ctypes.data, bad.size)])
    with pytest.raises(JpegGpuError):
        jb.parse_batch(items, threads=2)
    assert items[0].status == 0 and items[1].status != 0
    assert load_library().jpeggpu_b200_parse_batch(None, 0, 1, None) == 1  # INVALID_ARGUMENT
    assert load_library().jpeggpu_b200_decode_batch(None, 0, None, None, None, 1) == 1


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    files = _files()
    local = _parse_share(files, rank, world)
    everyone = jb.gather_counts(local, world)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        q.put(everyone)


def test_two_ranks_cover_the_batch_on_gloo():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    everyone = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    files = _files()
    assert [e["rank"] for e in everyone] == [0, 1]
    assert sum(e["items"] for e in everyone) == len(files)
    names = set()
    for e in everyone:
        assert not (names & set(e["sizes"]))
        names |= set(e["sizes"])
    assert names == {n for n, _ in files}
    single = _parse_share(files, 0, 1)
    for e in everyone:
        for k, v in e["sizes"].items():
            assert tuple(v) == tuple(single["sizes"][k])


@pytest.mark.gpu
@pytest.mark.parametrize("lanes", [1, 3])
def test_native_batch_decode_matches_oracle(lanes):
    import torch
    from jpeggpu_b200.torch_io import pinned_bytes
    files = _files() * 3  # 21 items of mixed geometry, more items than staging slots per lane
    bd = jb.BatchDecoder("cuda:0", lanes=lanes)
    bd.prepare([pinned_bytes(d) for _, d in files])
    bd.run()
    bd.synchronize()
    for i, (name, data) in enumerate(files):
        ref = ob.decode(data)
        assert bd.items[i].status == 0
        for c, p in enumerate(bd.planes[i]):
            assert np.array_equal(p.cpu().numpy(), ref["pixels"][c]), f"{name} item {i} comp {c}"
    torch.cuda.synchronize()


@pytest.mark.gpu
def test_one_decoder_many_images_in_flight():
    """The staging ring: one decoder object, 12 different images enqueued back to back on one
    stream without any host synchronisation in between."""
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    files = _files() + _files()[::-1]
    sess = DecodeSession("cuda:0")
    hosts = [pinned_bytes(d) for _, d in files]
    need = 0
    for h in hosts:
        sess.parse(h)
        need = max(need, sess.dec.get_buffer_size())
    sess.scratch = torch.empty(need, dtype=torch.uint8, device="cuda:0")
    outs = []
    for h in hosts:
        sess.parse(h)
        planes = sess.alloc_planes()
        sess.transfer()
        sess.decode(planes)
        outs.append((planes, sess.info))
    torch.cuda.synchronize()
    for (planes, info), (name, data) in zip(outs, files):
        ref = ob.decode(data)
        for c, p in enumerate(planes):
            assert np.array_equal(p[:, :info.sizes_x[c]].cpu().numpy(), ref["pixels"][c]), name


def test_batch_sizes_cover_every_file_once():
    """bench.py's batch leg: the sizes of a rank's fused batches (ramp at both ends) add up to the
    rank's files, none exceeds the batch size, and the shape follows the files per rank."""
    import argparse
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    for n in (0, 1, 7, 100, 511, 512, 513, 1024, 2048, 4096, 4097):
        for chunk in (8, 16, 32, 64, 128, 256):
            sizes = bench.batch_sizes(n, chunk)
            assert sum(sizes) == n and all(0 < x <= chunk for x in sizes), (n, chunk, sizes)
            if n >= 4 * chunk and chunk > 8:
                assert sizes[0] == max(8, chunk // 4) and sizes[-1] == sizes[0]
    shapes = [bench.batch_shape(argparse.Namespace(batch_images=4096, batch_chunk=0, batch_depth=0), w)
              for w in (1, 2, 4, 8)]
    assert shapes == [(128, 3), (128, 3), (64, 4), (32, 4)]
    assert bench.batch_shape(argparse.Namespace(batch_images=4096, batch_chunk=16, batch_depth=2), 8) == (16, 2)
