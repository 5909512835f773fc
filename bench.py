#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native baseline-JPEG decoder.

Workload (BASELINE.json configs[1]): one synthetic 4032x3024 4:2:0 q90 baseline JPEG written by the
reference's bundled stb_image_write (data/_gen/synth_4032x3024_s6510_q90.jpg, tools/gen_inputs.py).
A "step" decodes that image once.  Metric: MP/s = frame megapixels decoded per second.

Both arms print the same quantities (so that the driver's ratios compare like with like):
  value : DEVICE-RESIDENT throughput.  The compressed bytes + tables are already in HBM
          (jpeggpu_decoder_transfer done before the timed region); each step is one
          jpeggpu_decoder_decode, timed with CUDA events on the launching stream, L2 flushed
          between steps.
  e2e   : the reference's own benchmark protocol (benchmark/benchmark_jpeggpu.hpp:69-108) through
          the C ABI with HOST input: parse_header + get_buffer_size + transfer (H2D) + decode +
          a device->host read of a 4 KiB sample of the result + stream sync, wall clock.
  configs   : device / e2e milliseconds of the other single-image configurations of BASELINE.json
              (IMG_6510.JPG, 48 MP without restart markers, one 2 MP image), same two protocols.
  sustained : >= 1 s of back-to-back decodes of the workload image without L2 flush (clocks under
              sustained load).
  batch     : BASELINE.json configs[4]: 4096 distinct synthetic 1920x1080 files (seeds 0..4095),
              pre-loaded in pinned host memory, file i on rank i mod N; wall clock from the first
              enqueue to the last synchronisation, every byte uploaded inside the timed region;
              images/s = 4096 / max over ranks.  Ours: fused batches (K images per kernel launch,
              include/jpeggpu_b200.h); reference: its per-image caller loop.
  --impl reference : the UNMODIFIED reference decoder built from /root/reference/src into
          oracle/_ref/libjpeggpu_ref.so (its only implementation of the path is CUDA; it has no CPU
          decode).  Falls back to the CPU oracle port if that library is absent.

N > 1 (torchrun): every rank runs its own replica of the workload on its own GPU ("replicas only":
a single image does not shard); the batch is partitioned.  NO collective on the data path: the
process group only serves the barrier and the max over ranks of the timings.
"""
import argparse
import concurrent.futures
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

W, H, SEED, QUALITY = 4032, 3024, 6510, 90
MEGAPIXELS = W * H / 1e6
METRIC = "decode throughput, 12MP 4:2:0 baseline JPEG"
BATCH_W, BATCH_H = 1920, 1080


def load_workload():
    import gen_inputs
    p = gen_inputs.ensure_synth(W, H, SEED, QUALITY)
    if p is None:
        raise SystemExit("workload file missing and oracle/_ref/stbgen not built: run build() first")
    with open(p, "rb") as f:
        return f.read(), os.path.basename(p)


def workload_string(fname, nbytes):
    return f"{fname}: {W}x{H} 4:2:0 q90 stb_image_write, {nbytes} B"


def side_configs():
    """(key, bytes) of the other single-image configurations (those whose files are available)."""
    import gen_inputs
    out = []
    for key, path in (("IMG_6510.JPG 12MP 4:2:0 DRI=252", gen_inputs.ensure_img6510()),
                      ("stb 48MP 8064x6048 4:2:0 q90 no DRI", gen_inputs.ensure_synth(8064, 6048, 48)),
                      ("stb 2MP 1920x1080 4:2:0 q90", gen_inputs.ensure_synth(1920, 1080, 1))):
        if path and os.path.exists(path):
            out.append((key, open(path, "rb").read()))
    return out


def batch_paths(indices):
    """Paths of the batch files with the given seeds, generating the missing ones in parallel with
    the reference's bundled stb encoder (oracle/_ref/stbgen)."""
    import gen_inputs
    d = os.path.join(gen_inputs.GEN, f"batch_{BATCH_W}x{BATCH_H}")
    os.makedirs(d, exist_ok=True)
    paths = {i: os.path.join(d, f"s{i}.jpg") for i in indices}
    missing = [i for i in indices if not os.path.exists(paths[i])]
    if missing:
        if not os.path.exists(gen_inputs.STBGEN):
            return None

        def make(i):
            tmp = paths[i] + f".tmp{os.getpid()}"
            subprocess.run([gen_inputs.STBGEN, tmp, str(BATCH_W), str(BATCH_H), str(i), str(QUALITY), "1"],
                           check=True)
            os.replace(tmp, paths[i])

        workers = max(1, (os.cpu_count() or 8) // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))))
        with concurrent.futures.ThreadPoolExecutor(max_workers=workers) as ex:
            list(ex.map(make, missing))
    return [paths[i] for i in indices]


class ClockSampler:
    """nvidia-smi clock/throttle sampling during the timed regions (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, power, reasons = [], [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1]))
                    out["sm_max_mhz"] = float(f[2])
                    power.append(float(f[3]))
                except ValueError:
                    continue
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                    "sw_power_cap"), f[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            # "under load": the samples in the upper half of the power draw (the sampler also runs
            # through the idle gaps between the legs)
            cut = sorted(power)[len(power) // 2] if power else 0.0
            loaded = sorted(s for s, p in zip(sm, power) if p >= cut) or sorted(sm)
            out["sm_mhz"] = loaded[len(loaded) // 2]
            out["sm_mhz_min"] = min(sm)
            out["power_w_max"] = max(power) if power else None
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def dist_setup():
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group(backend="nccl", device_id=torch.device(f"cuda:{local}"))
    return rank, world, local


def barrier(world):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(value, world):
    import torch
    if world == 1:
        return value
    import torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def stage_table(kt_avg_ms, st):
    """Per-kernel algorithmic bytes and achieved GB/s."""
    Sb, Sd, N, D = st.scan_bytes, st.destuffed_bytes, st.num_subsequences, st.num_data_units
    Cc = 64 * D
    P = W * H + 2 * ((W + 1) // 2) * ((H + 1) // 2)
    bytes_of = {
        "destuff_kernel": Sb + Sd,
        "huff_sync": Sd + 20 * N + 2 * Cc + 2 * D,  # reads the scan, writes 16 B records + segment
                                                    # ids, zero-fills coefficients + DC array
        "huff_write": Sd + 24 * N + 2 * D,          # + the scattered non-zero coefficients
        "dc_scan": 4 * D,                           # compact DC array, in place
        "idct_kernel": 2 * Cc + 2 * D + P,
    }
    rows = {}
    for name, ms in kt_avg_ms.items():
        b = bytes_of.get(name, 0)
        rows[name] = {"ms": round(ms, 5), "bytes": b,
                      "gbs": round(b / (ms * 1e-3) / 1e9, 2) if ms > 0 else None}
    return rows


NCU_SUMMARY = "profiles/r2_ncu_full_summary.json"


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of `kernel` per launch from the committed
    `ncu --set full` capture of this workload on the current tree, else None."""
    try:
        rows = json.load(open(os.path.join(ROOT, NCU_SUMMARY)))
        for r in rows:
            if r["kernel"] == kernel:
                m = r["metrics"]
                scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

                def as_bytes(v):
                    num, _, unit = str(v).partition(" ")
                    return float(num) * scale.get(unit.strip(), 1e6)
                return int(as_bytes(m["dram__bytes_read.sum"]) + as_bytes(m["dram__bytes_write.sum"]))
    except Exception:
        pass
    return None


def cpu_baseline_port(data, budget_s=10.0):
    """The oracle (CPU restatement, 1 thread) decoding the same image, bounded to ~budget_s."""
    import oracle_bindings as ob
    t0 = time.perf_counter()
    r = ob.decode(data)
    assert r["status"] == 0
    first = time.perf_counter() - t0
    n = max(2, min(64, int(budget_s / max(first, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(n):
        ob.decode(data)
    dt = time.perf_counter() - t0
    return {"value": round(n * MEGAPIXELS / dt, 2), "unit": "MP/s", "cores": 1, "nproc": os.cpu_count(),
            "kind": "port",
            "sample": f"{n} full decodes of the workload image by oracle/jpeg_oracle.c (1 thread)"}


def reference_parse_us(data, iters=200):
    """Host-only marker parse of the reference (src/reader.cpp) via oracle/_ref, 1 thread."""
    try:
        import numpy as np
        import ref_bindings as rb
        if not rb.available():
            return None
        buf = np.frombuffer(data, dtype=np.uint8)
        rd = rb.RefDecoder()
        rd.parse(buf.ctypes.data, len(data))
        t0 = time.perf_counter()
        for _ in range(iters):
            rd.parse(buf.ctypes.data, len(data))
        dt = time.perf_counter() - t0
        rd.close()
        return round(dt / iters * 1e6, 2)
    except Exception:
        return None


def e2e_loop(step_fn, flush_fn, steps, warmup):
    """Wall-clock per-step timing with a sync on both sides; L2 flushed outside the timer."""
    import torch
    for _ in range(warmup):
        step_fn()
    times = []
    for _ in range(steps):
        flush_fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        step_fn()
        times.append(time.perf_counter() - t0)
    return times


def device_loop(decode_fn, flush_fn, steps, warmup):
    """CUDA events around decode_fn on the current stream, L2 flushed before every step; ms list."""
    import torch
    stream = torch.cuda.current_stream()
    for _ in range(warmup):
        decode_fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, b in ev:
        flush_fn()
        a.record(stream)
        decode_fn()
        b.record(stream)
    torch.cuda.synchronize()
    return [a.elapsed_time(b) for a, b in ev]


def median(xs):
    xs = sorted(xs)
    return xs[len(xs) // 2]


def nvjpeg_leg(host, nbytes, flush_fn, steps, warmup):
    """nvJPEG (local toolkit) on the same file and box: the decoupled OUTPUT_UNCHANGED protocol of
    the reference's own nvJPEG benchmark (benchmark/benchmark_nvjpeg.hpp:75-185) — stream parse +
    host phase + transfer + device phase + sync, wall clock, pinned input, L2 flushed between
    steps.  One entry per backend; a backend nvJPEG refuses is recorded as unavailable."""
    import torch
    try:
        import nvjpeg_bindings as nb
        if not nb.available():
            return {"unavailable": "oracle/_ref/libnvjpeg_harness.so not built"}
    except Exception as e:  # pragma: no cover
        return {"unavailable": repr(e)}
    out = {}
    s = torch.cuda.current_stream().cuda_stream
    for backend in ("gpu_hybrid", "hybrid", "hardware"):
        try:
            nj = nb.NvJpeg(backend)
        except Exception as e:
            out[backend] = {"unavailable": str(e)}
            continue
        try:
            ch, ws, hs = nj.info(host.data_ptr(), nbytes)
            planes = [torch.zeros((hs[c], ws[c]), dtype=torch.uint8, device="cuda") for c in range(ch)]
            ptrs = [p.data_ptr() for p in planes]
            pitches = [p.stride(0) for p in planes]

            def step():
                st = nj.decode(host.data_ptr(), nbytes, ptrs, pitches, s, True)
                if st:
                    raise RuntimeError(f"nvjpeg decode status {st}")

            n = min(steps, 20) if backend == "gpu_hybrid" else max(5, min(steps, 10))
            times = e2e_loop(step, flush_fn, n, warmup)
            total = sum(times)
            out[backend] = {"value": round(n * MEGAPIXELS / total, 1), "unit": "MP/s",
                            "ms_per_image": round(total / n * 1e3, 4), "steps": n}
        except Exception as e:
            out[backend] = {"unavailable": str(e)}
        finally:
            nj.close()
    out["protocol"] = ("nvjpegJpegStreamParse+DecodeJpegHost+TransferToDevice+DecodeJpegDevice+sync, "
                       "NVJPEG_OUTPUT_UNCHANGED, wall clock")
    return out


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------

def ours_single(sess, host, flush_l2, sample, steps, warmup):
    """(device ms list, e2e seconds list) of one file through DecodeSession `sess`."""
    import torch
    sess.parse(host)
    planes = sess.alloc_planes()
    sess.transfer()
    dev = device_loop(lambda: sess.decode(planes), flush_l2, steps, warmup)

    def e2e_step():
        sess.parse(host)
        sess.transfer()
        sess.decode(planes)
        sample.copy_(planes[0].view(-1)[:4096], non_blocking=True)
        torch.cuda.synchronize()

    e2e = e2e_loop(e2e_step, flush_l2, steps, warmup)
    return dev, e2e, planes


def ours_batch(rank, world, device, total_images, chunk, depth, repeats=3):
    """BASELINE config 5 on this rank: files i with i % world == rank, through fused batches."""
    import numpy as np
    import torch
    from jpeggpu_b200.batch import FusedBatch, make_items
    from jpeggpu_b200 import BatchItem
    mine = list(range(rank, total_images, world))
    paths = batch_paths(mine)
    if paths is None:
        return {"unavailable": "batch files missing and oracle/_ref/stbgen not built"}
    blobs = [open(p, "rb").read() for p in paths]
    # all files back to back in ONE pinned allocation (16-byte aligned starts)
    offs, at = [], 0
    for b in blobs:
        offs.append(at)
        at += (len(b) + 15) & ~15
    host = torch.empty(max(at, 16), dtype=torch.uint8, pin_memory=True)
    hv = host.numpy()
    for o, b in zip(offs, blobs):
        hv[o:o + len(b)] = np.frombuffer(b, dtype=np.uint8)
    items = make_items([(host.data_ptr() + o, len(b)) for o, b in zip(offs, blobs)])
    n = len(blobs)
    sizes = batch_sizes(n, chunk)
    chunks, at0 = [], 0
    for c in sizes:
        chunks.append((at0, c))
        at0 += c
    assert at0 == n
    lanes = []
    for _ in range(depth):
        fb = FusedBatch()
        fb.set_host_arena(host.data_ptr(), host.numel())
        lanes.append({"fb": fb, "stream": torch.cuda.Stream(device=device), "arena": None, "planes": None})
    isz = C_sizeof(BatchItem)
    # Output planes: one set per image of a chunk slot and lane, reused by the later chunks of the
    # lane (the reference benchmark reuses ONE output image; stream order keeps this safe).  The
    # caller's buffers are named in the items before the clock starts - all files share one geometry.
    ImgInfo0 = (BATCH_H, BATCH_W, (BATCH_H + 1) // 2, (BATCH_W + 1) // 2)
    for ln in lanes:
        ln["planes"] = [[torch.empty((ImgInfo0[0], ImgInfo0[1]), dtype=torch.uint8, device=device),
                         torch.empty((ImgInfo0[2], ImgInfo0[3]), dtype=torch.uint8, device=device),
                         torch.empty((ImgInfo0[2], ImgInfo0[3]), dtype=torch.uint8, device=device)]
                        for _ in range(chunk)]
    for k, (first, count) in enumerate(chunks):
        ln = lanes[k % depth]
        for j in range(count):
            it = items[first + j]
            for c, p in enumerate(ln["planes"][j]):
                it.img.image[c] = p.data_ptr()
                it.img.pitch[c] = p.stride(0)

    # One host thread per batch in flight: thread w plans and enqueues chunks w, w + depth, ... on
    # its own batch object and stream (the C calls release the GIL, so the threads really overlap:
    # planning is ~9 us per file on one thread, the device needs ~12).
    def lane_worker(li, which):
        torch.cuda.set_device(device)
        ln = lanes[li]
        plan_s = enq_s = 0.0
        for k in range(li, len(which), depth):
            first, count = which[k]
            sub = C_slice(items, first, isz)
            h0 = time.perf_counter()
            need, _ = ln["fb"].plan(sub, count=count)
            h1 = time.perf_counter()
            if ln["arena"].numel() < need:
                # (never with one geometry: the arenas are sized for the largest batch up front.  A
                # new arena must not replace one that batches in flight still use.)
                ln["stream"].synchronize()
                ln["arena"] = torch.empty(need + need // 8, dtype=torch.uint8, device=device)
            ln["fb"].decode(sub, ln["arena"].data_ptr(), ln["arena"].numel(), ln["stream"].cuda_stream,
                            count=count)
            plan_s += h1 - h0
            enq_s += time.perf_counter() - h1
        ln["stream"].synchronize()
        return plan_s, enq_s

    pool = concurrent.futures.ThreadPoolExecutor(max_workers=depth)
    # every lane's arena, sized for the largest batch before anything runs
    largest = max(chunks, key=lambda fc: fc[1])
    for ln in lanes:
        need, _ = ln["fb"].plan(C_slice(items, largest[0], isz), count=largest[1])
        ln["arena"] = torch.empty(need + need // 8, dtype=torch.uint8, device=device)

    def run(which, timed):
        t0 = time.perf_counter()
        res = list(pool.map(lambda li: lane_worker(li, which), range(depth)))
        dt = time.perf_counter() - t0
        return dt, max(r[0] for r in res), max(r[1] for r in res)

    # warm-up: one untimed pass over the whole batch (allocations, pinned staging, kernel attributes,
    # the threads' CUDA contexts), then `repeats` timed passes - every pass uploads every file again;
    # the median pass is reported, all of them are listed
    run(chunks, False)
    passes = []
    for _ in range(max(1, repeats)):
        barrier(world)
        passes.append(run(chunks, True))
    barrier(world)
    dt, plan_s, enq_s = sorted(passes, key=lambda r: r[0])[len(passes) // 2]
    host_s = plan_s + enq_s
    info0 = items[0].info
    assert (info0.sizes_y[0], info0.sizes_x[0], info0.sizes_y[1], info0.sizes_x[1]) == ImgInfo0
    # spot check: the last image of the last chunk against the per-image path
    ok = None
    try:
        from jpeggpu_b200.torch_io import decode_file
        first, count = chunks[-1]
        ln = lanes[(len(chunks) - 1) % depth]
        info, ref_planes, _ = decode_file(blobs[first + count - 1], device=str(device))
        ok = all(np.array_equal(ln["planes"][count - 1][c].cpu().numpy(), ref_planes[c]) for c in range(len(ref_planes)))
    except Exception as e:  # pragma: no cover
        ok = f"check failed: {e!r}"
    pool.shutdown()
    for ln in lanes:
        ln["fb"].close()
    total_bytes = sum(len(b) for b in blobs)
    return {"seconds": dt, "images": n, "host_enqueue_seconds": host_s, "plan_seconds": plan_s,
            "bytes": total_bytes, "matches_per_image_path": ok, "seconds_all_passes": [round(r[0], 5) for r in passes]}


def batch_sizes(n, chunk):
    """Sizes of the fused batches of one rank's n files: `chunk` images each, with a short ramp at
    both ends (a quarter of `chunk`, at least 8, doubling) - the first kernels start after planning
    and uploading a few files instead of `chunk`, and the last batch the pipeline has to drain is
    a small one."""
    if chunk <= 8 or n < 4 * chunk:
        return [chunk] * (n // chunk) + ([n % chunk] if n % chunk else [])
    head, c = [], max(8, chunk // 4)
    while c < chunk:
        head.append(c)
        c *= 2
    tail = head[::-1]
    body = n - sum(head) - sum(tail)
    return head + [chunk] * (body // chunk) + ([body % chunk] if body % chunk else []) + tail


def batch_shape(args, world):
    """Images per fused batch and batches in flight.  Large batches amortise launches and the
    descriptor upload; small ones shorten the fill and drain of the pipeline, which is what counts
    when a rank has few images (ours_batch also ramps the batch size at both ends).  Measured on
    B200s, 4096 files: 1 GPU 128 x 3 -> 88 k images/s, 64 x 4 -> 87 k; 512 files on one GPU (a
    rank's share on 8): 64 x 4 -> 81.4 k, 32 x 4 -> 80.0 k, 16 x 4 -> 75.9 k; on 8 GPUs, where eight
    processes share 32 host cores, 32 x 4 -> 575-585 k, 64 x 4 -> 550-562 k, 128 x 3 -> 490-510 k
    (profiles/r2_final_batch_scale_*.txt)."""
    per_rank = max(1, args.batch_images // world)
    chunk = args.batch_chunk
    if chunk <= 0:
        chunk = 128 if per_rank >= 2048 else 64 if per_rank >= 1024 else 32
    depth = args.batch_depth if args.batch_depth > 0 else (3 if chunk >= 128 else 4)
    return chunk, depth


def C_sizeof(t):
    import ctypes
    return ctypes.sizeof(t)


def C_slice(arr, first, item_size):
    """Pointer to element `first` of a ctypes array (no copy)."""
    import ctypes
    return ctypes.cast(ctypes.addressof(arr) + first * item_size, ctypes.c_void_p)


def run_ours(args):
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    rank, world, local = dist_setup()
    device = torch.device(f"cuda:{local}")
    data, fname = load_workload()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()  # before any warm-up: the legs are short
    sess = DecodeSession(f"cuda:{local}")
    host = pinned_bytes(data)
    info = sess.parse(host)
    assert (info.sizes_x[0], info.sizes_y[0]) == (W, H)
    st = sess.dec.scan_stats(0)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=sess.device)  # > 126 MB L2
    sample = torch.empty(4096, dtype=torch.uint8, pin_memory=True)
    stream = torch.cuda.current_stream()

    def flush_l2():
        flush.fill_(1)

    # ---- device-resident throughput ("value") and end to end ("e2e") --------------------------
    sess.parse(host)
    planes = sess.alloc_planes()
    sess.transfer()
    for _ in range(args.warmup):
        sess.decode(planes)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    barrier(world)
    for a, b in ev:
        flush_l2()
        a.record(stream)
        sess.decode(planes)
        b.record(stream)
    barrier(world)
    dev_ms = max_over_ranks(sum(a.elapsed_time(b) for a, b in ev), world)

    def e2e_step():
        sess.parse(host)
        sess.transfer()
        sess.decode(planes)
        sample.copy_(planes[0].view(-1)[:4096], non_blocking=True)
        torch.cuda.synchronize()

    barrier(world)
    e2e_times = e2e_loop(e2e_step, flush_l2, args.steps, args.warmup)
    barrier(world)
    e2e_s = max_over_ranks(sum(e2e_times), world)
    h2d = sess.dec.transfer_size()

    # ---- sustained: >= 1 s of back-to-back decodes, inputs resident, no flush -----------------
    n_sus = max(200, int(1.2 / max(dev_ms / args.steps * 1e-3, 1e-6)))
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier(world)
    s0.record(stream)
    for _ in range(n_sus):
        sess.decode(planes)
    s1.record(stream)
    barrier(world)
    sus_ms = max_over_ranks(s0.elapsed_time(s1), world)

    # ---- pipelined throughput: several images in flight on one GPU (extra) ---------------------
    depth = 4
    lanes = []
    for _ in range(depth):
        ls = DecodeSession(f"cuda:{local}")
        ls.parse(host)
        lp = ls.alloc_planes()
        lst = torch.cuda.Stream()
        ls.transfer(lst)
        lanes.append((ls, lp, lst))
    torch.cuda.synchronize()
    rounds = max(4, args.steps // depth)
    for ls, lp, lst in lanes:
        ls.decode(lp, lst)
    torch.cuda.synchronize()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier(world)
    p0.record(stream)
    for ls, lp, lst in lanes:
        lst.wait_event(p0)
    for _ in range(rounds):
        for ls, lp, lst in lanes:
            ls.decode(lp, lst)
    for ls, lp, lst in lanes:
        stream.wait_stream(lst)
    p1.record(stream)
    barrier(world)
    pipe_ms = max_over_ranks(p0.elapsed_time(p1), world)
    pipe_value = world * rounds * depth * MEGAPIXELS / (pipe_ms * 1e-3)
    for ls, lp, lst in lanes:
        ls.close()
    del lanes

    # ---- per-kernel times (profiling pass, separate from the timed regions) -------------------
    sess.dec.set_profiling(True)
    acc = {}
    reps = max(3, min(args.steps, 20))
    for _ in range(reps):
        flush_l2()
        sess.decode(planes)
        torch.cuda.synchronize()
        for k, (ms, n) in sess.dec.kernel_times().items():
            acc[k] = acc.get(k, 0.0) + ms
    sess.dec.set_profiling(False)
    kt = {k: v / reps for k, v in acc.items()}
    launches = st.kernel_launches

    # ---- optional RGB output (extra) ------------------------------------------
    rgb = torch.empty((H, W, 3), dtype=torch.uint8, device=sess.device)
    rgb_ms = median(device_loop(lambda: sess.decode_rgbi(rgb), flush_l2, reps, 3))
    del rgb

    # ---- the other single-image configurations (rank 0; not part of the headline) -------------
    configs = {}
    if rank == 0 and not args.no_configs:
        for key, blob in side_configs():
            cs = DecodeSession(f"cuda:{local}")
            ch = pinned_bytes(blob)
            n_it = max(5, min(args.steps, 20))
            d, e, pl = ours_single(cs, ch, flush_l2, sample, n_it, 3)
            ci = cs.info
            configs[key] = {"device_ms": round(median(d), 5), "e2e_ms": round(median(e) * 1e3, 5),
                            "megapixels": round(ci.sizes_x[0] * ci.sizes_y[0] / 1e6, 2), "bytes": len(blob),
                            "steps": n_it}
            del pl
            cs.close()

    # ---- batch (BASELINE config 5) ----------------------------------------------------------------
    batch = None
    if args.batch_images > 0:
        del flush
        torch.cuda.empty_cache()
        b_chunk, b_depth = batch_shape(args, world)
        r = ours_batch(rank, world, device, args.batch_images, b_chunk, b_depth, args.batch_repeats)
        if "seconds" in r:
            secs = max_over_ranks(r["seconds"], world)
            host_s = max_over_ranks(r["host_enqueue_seconds"], world)
            batch = {"images_per_s": round(args.batch_images / secs, 1), "images": args.batch_images,
                     "seconds": round(secs, 4), "unit": "images/s",
                     "megapixels_per_s": round(args.batch_images * BATCH_W * BATCH_H / 1e6 / secs, 1),
                     "workload": f"{args.batch_images} distinct stb {BATCH_W}x{BATCH_H} 4:2:0 q90 files (seeds 0..{args.batch_images - 1}), "
                                 f"pinned host memory, file i on rank i mod {world}",
                     "path": "jpeggpu_b200_batch_plan + jpeggpu_b200_batch_decode (fused: "
                             f"{b_chunk} images per kernel launch, {b_depth} batches in flight per GPU)",
                     "timing": "wall clock, first enqueue to last stream sync, H2D of every file inside; max over ranks",
                     "host_threads_per_gpu": b_depth,
                     "passes": f"1 untimed + {args.batch_repeats} timed passes over the whole batch, median reported",
                     "seconds_all_passes_rank0": r["seconds_all_passes"],
                     "host_busy_seconds_slowest_thread": round(host_s, 4),
                     "host_plan_seconds_rank0": round(r["plan_seconds"], 4),
                     "h2d_bytes_rank0": r["bytes"], "nproc": os.cpu_count(),
                     "matches_per_image_path": r["matches_per_image_path"]}
        else:
            batch = r

    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback 6650"
    rows = stage_table(kt, st)
    top = max(rows, key=lambda k: rows[k]["ms"])
    achieved = rows[top]["gbs"] or 0.0
    ms_per_step = dev_ms / args.steps
    value = world * MEGAPIXELS / (ms_per_step * 1e-3)
    e2e_value = world * args.steps * MEGAPIXELS / e2e_s
    cpu = cpu_baseline_port(data)
    rp = reference_parse_us(data)
    if rp is not None:
        cpu["reference_parse_us"] = rp
        cpu["reference_parse_note"] = "reference reader.cpp parse_header of the workload, 1 thread"
    out = {
        "metric": METRIC, "value": round(value, 1),
        "unit": "MP/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_per_step, 5), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int16/int32 fixed point", "data": "synthetic",
        "images_per_s": round(value / MEGAPIXELS, 1),
        "value_protocol": "device resident: CUDA events around jpeggpu_decoder_decode, L2 flushed",
        "config": {"workload": workload_string(fname, len(data)),
                   "scan": f"{st.num_segments} segment(s), {st.num_subsequences} subsequences of "
                           f"{st.subsequence_bytes} B",
                   "l2": "flushed (512 MiB fill) before every timed step",
                   "parallelism": f"replicas x{world}",
                   "collective": "none on the data path (the process group serves the barrier and "
                                 "the max over ranks of the timings only)"},
        "e2e": {"value": round(e2e_value, 1), "unit": "MP/s",
                "ms_per_image": round(e2e_s / args.steps * 1e3, 5),
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4096,
                "protocol": "parse_header+get_buffer_size+transfer+decode+4KiB D2H+sync, wall clock"},
        "sustained": {"value": round(world * n_sus * MEGAPIXELS / (sus_ms * 1e-3), 1), "unit": "MP/s",
                      "seconds": round(sus_ms * 1e-3, 3), "decodes": n_sus,
                      "note": "back-to-back decodes of the workload image, inputs resident, no L2 flush"},
        "throughput_pipelined": {"value": round(pipe_value, 1), "unit": "MP/s", "images_in_flight": depth,
                                 "images": rounds * depth * world,
                                 "note": "same image, 4 decoders/streams per GPU, inputs resident, "
                                         "CUDA events; L2 not flushed between images"},
        "rgb_output": {"value": round(MEGAPIXELS / (rgb_ms * 1e-3), 1), "unit": "MP/s",
                       "ms_per_image": round(rgb_ms, 5), "steps": reps,
                       "note": "jpeggpu_b200_decode_rgbi: same scan kernels, IDCT fused with chroma "
                               "replication + YCbCr->RGB (bit-exact vs the reference's conv_to_rgbi); "
                               "inputs resident, CUDA events, L2 flushed; one GPU"},
        "configs": configs,
        "batch": batch,
        "gpu_launches": launches * args.steps,
        "launches_per_image": launches,
        "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": round(achieved / peak, 5), "traffic": ncu_traffic(top),
                     "peak_source": peak_src,
                     "traffic_source": NCU_SUMMARY + " (ncu --set full, cold caches; the zero-fill stays in L2)",
                     "note": "Huffman kernels are issue/latency bound (bit-serial); see stages"},
        "stages": rows,
        "sum_kernel_ms": round(sum(v["ms"] for v in rows.values()), 5),
        "cpu_baseline": cpu,
        "clocks": clocks,
    }
    emit(out)


# ---------------------------------------------------------------------------------------------
# reference arm
# ---------------------------------------------------------------------------------------------

class RefSession:
    """The reference decoder (oracle/_ref/libjpeggpu_ref.so) on one file: same two protocols."""

    def __init__(self, blob):
        import numpy as np
        import torch
        import ref_bindings as rb
        self.torch = torch
        self.host = torch.empty(len(blob), dtype=torch.uint8, pin_memory=True)
        self.host.numpy()[:] = np.frombuffer(blob, dtype=np.uint8)
        self.n = len(blob)
        self.rd = rb.RefDecoder()
        st, self.info = self.rd.parse(self.host.data_ptr(), self.n)
        if st != 0:
            raise RuntimeError(f"reference parse status {st}")
        self.planes = [torch.zeros((self.info.sizes_y[c], self.info.sizes_x[c]), dtype=torch.uint8, device="cuda")
                       for c in range(self.info.num_components)]
        self.need = self.rd.buffer_size()
        self.tmp = torch.empty(self.need, dtype=torch.uint8, device="cuda")
        self.ptrs = [p.data_ptr() for p in self.planes]
        self.pitches = [p.stride(0) for p in self.planes]
        self.sample = torch.empty(4096, dtype=torch.uint8, pin_memory=True)

    def stream(self):
        return self.torch.cuda.current_stream().cuda_stream

    def e2e_step(self):
        st, _ = self.rd.parse(self.host.data_ptr(), self.n)
        n = self.rd.buffer_size()
        assert st == 0 and n <= self.need
        s = self.stream()
        assert self.rd.transfer(self.tmp.data_ptr(), self.need, s) == 0
        assert self.rd.decode(self.ptrs, self.pitches, self.tmp.data_ptr(), self.need, s) == 0
        self.sample.copy_(self.planes[0].view(-1)[:4096], non_blocking=True)
        self.torch.cuda.synchronize()

    def decode_only(self):
        assert self.rd.decode(self.ptrs, self.pitches, self.tmp.data_ptr(), self.need, self.stream()) == 0

    def close(self):
        self.rd.close()


def reference_batch(rank, world, total_images):
    """The reference's caller loop (benchmark/benchmark_jpeggpu.hpp:69-89) over this rank's files:
    parse + size + transfer + decode per image on one stream, one sync at the end."""
    import numpy as np
    import torch
    import ref_bindings as rb
    mine = list(range(rank, total_images, world))
    paths = batch_paths(mine)
    if paths is None:
        return None
    blobs = [open(p, "rb").read() for p in paths]
    hosts = []
    for b in blobs:
        t = torch.empty(len(b), dtype=torch.uint8, pin_memory=True)
        t.numpy()[:] = np.frombuffer(b, dtype=np.uint8)
        hosts.append(t)
    rd = rb.RefDecoder()
    st, info = rd.parse(hosts[0].data_ptr(), hosts[0].numel())
    planes = [torch.zeros((info.sizes_y[c], info.sizes_x[c]), dtype=torch.uint8, device="cuda")
              for c in range(info.num_components)]
    need = rd.buffer_size()
    tmp = torch.empty(need + need // 4, dtype=torch.uint8, device="cuda")
    ptrs, pitches = [p.data_ptr() for p in planes], [p.stride(0) for p in planes]
    s = torch.cuda.current_stream().cuda_stream

    def run(hs):
        t0 = time.perf_counter()
        for h in hs:
            st, _ = rd.parse(h.data_ptr(), h.numel())
            assert st == 0 and rd.buffer_size() <= tmp.numel()
            # the reference keeps pointers into ONE pinned table block per decoder: its benchmark
            # synchronises after every image (benchmark_jpeggpu.hpp:88), and so must this loop
            assert rd.transfer(tmp.data_ptr(), tmp.numel(), s) == 0
            assert rd.decode(ptrs, pitches, tmp.data_ptr(), tmp.numel(), s) == 0
            torch.cuda.synchronize()
        return time.perf_counter() - t0

    run(hosts[:4])
    barrier(world)
    dt = run(hosts)
    barrier(world)
    rd.close()
    return {"seconds": dt, "images": len(hosts)}


def run_reference(args):
    """The reference arm: the UNMODIFIED reference decoder (oracle/_ref/libjpeggpu_ref.so) driven
    through its own public API, one replica per rank; CPU oracle port if it cannot run."""
    data, fname = load_workload()
    base = {"impl": "reference", "metric": METRIC,
            "unit": "MP/s", "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "data": "synthetic",
            "config": {"workload": workload_string(fname, len(data))}}
    gpu_ok = False
    try:
        import torch
        import ref_bindings as rb
        gpu_ok = torch.cuda.is_available() and rb.available()
    except Exception:
        gpu_ok = False
    if not gpu_ok:
        if int(os.environ.get("RANK", "0")) != 0:
            return
        cpu = cpu_baseline_port(data, budget_s=20.0)
        base.update({"n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "value": cpu["value"],
                     "ms_per_step": round(MEGAPIXELS / cpu["value"] * 1e3, 3),
                     "cpu_baseline": cpu, "dtype": "int16/int32 fixed point",
                     "e2e": {"value": cpu["value"], "unit": "MP/s", "h2d_bytes_per_step": 0,
                             "d2h_bytes_per_step": 0},
                     "note": "reference CUDA library unavailable: timed the CPU oracle port"})
        emit(base)
        return
    import torch
    rank, world, local = dist_setup()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")

    def flush_l2():
        flush.fill_(1)

    rs = RefSession(data)
    barrier(world)
    e2e_times = e2e_loop(rs.e2e_step, flush_l2, args.steps, args.warmup)
    barrier(world)
    e2e_s = max_over_ranks(sum(e2e_times), world)
    barrier(world)
    dev = device_loop(rs.decode_only, flush_l2, args.steps, args.warmup)
    barrier(world)
    dev_ms = max_over_ranks(sum(dev), world)
    n_sus = max(50, int(1.2 / max(dev_ms / args.steps * 1e-3, 1e-6)))
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier(world)
    s0.record()
    for _ in range(n_sus):
        rs.decode_only()
    s1.record()
    barrier(world)
    sus_ms = max_over_ranks(s0.elapsed_time(s1), world)
    configs = {}
    if rank == 0 and not args.no_configs:
        for key, blob in side_configs():
            try:
                cs = RefSession(blob)
                n_it = max(5, min(args.steps, 20))
                e = e2e_loop(cs.e2e_step, flush_l2, n_it, 3)
                d = device_loop(cs.decode_only, flush_l2, n_it, 3)
                configs[key] = {"device_ms": round(median(d), 5), "e2e_ms": round(median(e) * 1e3, 5),
                                "megapixels": round(cs.info.sizes_x[0] * cs.info.sizes_y[0] / 1e6, 2),
                                "bytes": len(blob), "steps": n_it}
                cs.close()
                del cs
            except Exception as ex:  # pragma: no cover
                configs[key] = {"unavailable": repr(ex)}
    nvj = nvjpeg_leg(rs.host, len(data), flush_l2, args.steps, args.warmup) if rank == 0 and not args.no_configs else None
    rs.close()
    batch = None
    if args.batch_images > 0:
        del flush
        torch.cuda.empty_cache()
        n_img = min(args.batch_images, args.reference_batch_images)
        r = reference_batch(rank, world, n_img)
        if r is not None:
            secs = max_over_ranks(r["seconds"], world)
            batch = {"images_per_s": round(n_img / secs, 1), "images": n_img, "seconds": round(secs, 4),
                     "unit": "images/s",
                     "megapixels_per_s": round(n_img * BATCH_W * BATCH_H / 1e6 / secs, 1),
                     "workload": f"the first {n_img} of the {args.batch_images} distinct stb {BATCH_W}x{BATCH_H} files, "
                                 f"pinned host memory, file i on rank i mod {world}",
                     "path": "the reference's caller loop (benchmark/benchmark_jpeggpu.hpp:69-89): parse_header + "
                             "get_buffer_size + transfer + decode + sync per image, one decoder per GPU",
                     "timing": "wall clock, first call to last sync; max over ranks", "nproc": os.cpu_count()}
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return
    ms_per_step = dev_ms / args.steps
    value = world * MEGAPIXELS / (ms_per_step * 1e-3)
    e2e_value = world * args.steps * MEGAPIXELS / e2e_s
    cpu = {"value": round(e2e_value, 1), "unit": "MP/s", "cores": 1, "nproc": os.cpu_count(), "kind": "reference",
           "sample": "the reference has no CPU decode path: this line times its CUDA decoder "
                     "(oracle/_ref/libjpeggpu_ref.so, built from /root/reference/src for sm_100) "
                     "driven by 1 host thread per GPU; value = its end-to-end rate (host-side parse inside)"}
    rp = reference_parse_us(data)
    if rp is not None:
        cpu["reference_parse_us"] = rp
    base.update({"n_gpus": world, "value": round(value, 1), "ms_per_step": round(ms_per_step, 5),
                 "images_per_s": round(value / MEGAPIXELS, 1),
                 "value_protocol": "device resident: CUDA events around jpeggpu_decoder_decode, L2 flushed",
                 "dtype": "int16/int32 fixed point",
                 "e2e": {"value": round(e2e_value, 1), "unit": "MP/s",
                         "ms_per_image": round(e2e_s / args.steps * 1e3, 5),
                         "h2d_bytes_per_step": len(data) + 4 * 64 + 4 * 7168,
                         "d2h_bytes_per_step": 4096,
                         "protocol": "parse_header+get_buffer_size+transfer+decode+4KiB D2H+sync, wall clock"},
                 "sustained": {"value": round(world * n_sus * MEGAPIXELS / (sus_ms * 1e-3), 1), "unit": "MP/s",
                               "seconds": round(sus_ms * 1e-3, 3), "decodes": n_sus},
                 "configs": configs, "batch": batch,
                 "cpu_baseline": cpu, "nvjpeg": nvj, "clocks": clocks})
    emit(base)


_RESULT_FD = None


def emit(obj):
    """The ONE JSON line of the contract, written to the process's original stdout."""
    line = json.dumps(obj) + "\n"
    if _RESULT_FD is None:
        sys.stdout.write(line)
        sys.stdout.flush()
    else:
        os.write(_RESULT_FD, line.encode())


def main():
    # Libraries talk on stdout (NCCL prints its version line there when NCCL_DEBUG is set in the
    # environment): keep the original stdout for the result line and point fd 1 at stderr.
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch-images", type=int, default=4096, help="BASELINE config 5 (0 disables the leg)")
    ap.add_argument("--batch-chunk", type=int, default=0,
                    help="images per fused batch (0: from the images per rank, see batch_shape)")
    ap.add_argument("--batch-depth", type=int, default=0, help="fused batches in flight per GPU (0: auto)")
    ap.add_argument("--batch-repeats", type=int, default=3, help="timed passes over the batch (median reported)")
    ap.add_argument("--reference-batch-images", type=int, default=1024,
                    help="images of the batch the reference arm's per-image loop decodes (bounded sample)")
    ap.add_argument("--no-configs", action="store_true", help="skip the side configurations and nvJPEG")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
