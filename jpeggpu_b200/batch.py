"""Batches of independent images (BASELINE config 5).

A single image does not shard across GPUs, a batch does trivially: rank r of a world of W
processes (one per GPU) owns items r, r + W, r + 2W, ... and decodes them on its own GPU through
jpeggpu_b200_decode_batch (native host threads, one per stream).  There is NO collective on the
data path; `gather_counts` (used by tools/batch_decode.py and the tests) only collects per-rank
bookkeeping.  Nothing here computes on the CPU: parsing and decoding happen inside the library.
"""
import ctypes as C

from . import BatchItem, JpegGpuError, SUCCESS, load_library


def partition(num_items: int, rank: int, world: int):
    """Indices of the batch items rank `rank` owns (round robin: equal shares +-1 and, for
    batches sorted by size, equal work)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return list(range(rank, num_items, world))


def make_items(host_buffers):
    """host_buffers: list of (address, size) of JPEG files in (pinned) host memory."""
    arr = (BatchItem * len(host_buffers))()
    for it, (ptr, size) in zip(arr, host_buffers):
        it.data = ptr
        it.size = size
    return arr


def parse_batch(items, threads=8, lib=None):
    """Host-only header parse of every item; returns the largest scratch requirement."""
    lib = lib or load_library()
    need = C.c_size_t()
    st = lib.jpeggpu_b200_parse_batch(items, len(items), threads, C.byref(need))
    if st != SUCCESS:
        raise JpegGpuError(st, "jpeggpu_b200_parse_batch")
    return need.value


def decode_batch(items, streams, scratch_ptrs, scratch_sizes, lib=None):
    """Enqueue the decode of every item; streams / scratch_*: one per lane (integers)."""
    lib = lib or load_library()
    n = len(streams)
    s = (C.c_void_p * n)(*streams)
    p = (C.c_void_p * n)(*scratch_ptrs)
    z = (C.c_size_t * n)(*scratch_sizes)
    st = lib.jpeggpu_b200_decode_batch(items, len(items), s, p, z, n)
    if st != SUCCESS:
        raise JpegGpuError(st, "jpeggpu_b200_decode_batch")


def gather_counts(local: dict, world: int):
    """All-gather a small dict of per-rank numbers (bookkeeping only, works on gloo and nccl)."""
    if world == 1:
        return [local]
    import torch.distributed as dist
    out = [None] * world
    dist.all_gather_object(out, local)
    return out


class BatchDecoder:
    """torch-side plumbing for one GPU: pinned inputs, `lanes` streams with one scratch each,
    one set of output planes per item (uniform batches may share a ring of plane sets)."""

    def __init__(self, device, lanes=4):
        import torch
        self.torch = torch
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        self.lanes = lanes
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(lanes)]
        self.scratch = []
        self.lib = load_library()

    def prepare(self, files, plane_sets=None):
        """files: list of pinned uint8 tensors.  Parses all headers (host threads), allocates
        scratch per lane and output planes.  plane_sets: number of distinct output plane sets
        (None = one per item); item i writes set i % plane_sets — with fewer sets than items,
        consecutive items of a lane overwrite each other's output in stream order, which is the
        reference benchmark's behaviour (one output image reused for every iteration)."""
        torch = self.torch
        self.files = files
        self.items = make_items([(f.data_ptr(), f.numel()) for f in files])
        need = parse_batch(self.items, threads=min(16, max(1, len(files))), lib=self.lib)
        if not self.scratch or self.scratch[0].numel() < need:
            self.scratch = [torch.empty(need, dtype=torch.uint8, device=self.device)
                            for _ in range(self.lanes)]
        n_sets = len(files) if plane_sets is None else min(plane_sets, len(files))
        self.planes = []
        for k in range(n_sets):
            info = self.items[k].info
            self.planes.append([torch.zeros((info.sizes_y[c], info.sizes_x[c]), dtype=torch.uint8,
                                            device=self.device) for c in range(info.num_components)])
        for i, it in enumerate(self.items):
            ps = self.planes[i % n_sets]
            if len(ps) != it.info.num_components or any(
                    (p.shape[1], p.shape[0]) != (it.info.sizes_x[c], it.info.sizes_y[c])
                    for c, p in enumerate(ps)):
                raise ValueError("plane_sets < items needs a batch of uniform geometry")
            for c, p in enumerate(ps):
                it.img.image[c] = p.data_ptr()
                it.img.pitch[c] = p.stride(0)
        return need

    def run(self):
        """Enqueue the whole batch (returns before the device is done)."""
        decode_batch(self.items, [s.cuda_stream for s in self.streams],
                     [t.data_ptr() for t in self.scratch], [t.numel() for t in self.scratch],
                     lib=self.lib)

    def synchronize(self):
        for s in self.streams:
            s.synchronize()


class FusedBatch:
    """One jpeggpu_b200_batch_t (include/jpeggpu_b200.h "fused batches"): K images per kernel
    launch.  plan() is host-only and returns the device arena size; decode() enqueues one upload of
    descriptors + tables, one memset, the file bytes (one copy when the files are contiguous in
    host memory) and five kernel launches on the caller's stream."""

    def __init__(self, lib=None):
        self.lib = lib or load_library()
        self.handle = C.c_void_p()
        st = self.lib.jpeggpu_b200_batch_create(C.byref(self.handle))
        if st != SUCCESS:
            raise JpegGpuError(st, "jpeggpu_b200_batch_create")

    def set_host_arena(self, base, nbytes):
        """Promise that [base, base + nbytes) is one host allocation holding the files: batches
        whose files lie close together inside it are uploaded with a single copy."""
        st = self.lib.jpeggpu_b200_batch_set_host_arena(self.handle, base, nbytes)
        if st != SUCCESS:
            raise JpegGpuError(st, "jpeggpu_b200_batch_set_host_arena")

    def plan(self, items, count=None, strict=True):
        """items: BatchItem array (or a slice made with ctypes pointer arithmetic); returns
        (arena bytes, status of the first item the fused path cannot take)."""
        n = len(items) if count is None else count
        need = C.c_size_t()
        st = self.lib.jpeggpu_b200_batch_plan(self.handle, items, n, C.byref(need))
        if strict and st != SUCCESS:
            raise JpegGpuError(st, "jpeggpu_b200_batch_plan")
        return need.value, st

    def decode(self, items, d_arena, arena_size, stream=0, count=None, strict=True):
        n = len(items) if count is None else count
        st = self.lib.jpeggpu_b200_batch_decode(self.handle, items, n, d_arena, arena_size, stream)
        if strict and st != SUCCESS:
            raise JpegGpuError(st, "jpeggpu_b200_batch_decode")
        return st

    def read_scan_info(self, index, d_arena, stream=0):
        from . import ScanInfo
        si = ScanInfo()
        st = self.lib.jpeggpu_b200_batch_read_scan_info(self.handle, index, d_arena, stream, C.byref(si))
        if st != SUCCESS:
            raise JpegGpuError(st, "jpeggpu_b200_batch_read_scan_info")
        return si

    def close(self):
        if self.handle:
            self.lib.jpeggpu_b200_batch_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
