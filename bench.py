#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native baseline-JPEG decoder.

Workload (BASELINE.json configs[1]): one synthetic 4032x3024 4:2:0 q90 baseline JPEG written by the
reference's bundled stb_image_write (data/_gen/synth_4032x3024_s6510_q90.jpg, tools/gen_inputs.py).
A "step" decodes that image once.  Metric: MP/s = frame megapixels decoded per second.

  value : device-resident throughput.  The compressed bytes + tables are already in HBM
          (jpeggpu_decoder_transfer done before the timed region); each step is one
          jpeggpu_decoder_decode, timed with CUDA events on the launching stream, L2 flushed
          between steps.
  e2e   : the reference's own benchmark protocol (benchmark/benchmark_jpeggpu.hpp:69-108) through
          the C ABI with HOST input: parse_header + get_buffer_size + transfer (H2D) + decode +
          a device->host read of a 4 KiB sample of the result + stream sync, wall clock.
  --impl reference : the same e2e protocol driven through the UNMODIFIED reference decoder built
          from /root/reference/src into oracle/_ref/libjpeggpu_ref.so (its only implementation of
          the path is CUDA; it has no CPU decode).  Falls back to the CPU oracle port if that
          library is absent.

N > 1 (torchrun): every rank decodes its own replica of the workload on its own GPU ("replicas
only": a single image does not shard; no collective on the data path), value = total MP / max time.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

W, H, SEED, QUALITY = 4032, 3024, 6510, 90
MEGAPIXELS = W * H / 1e6


def load_workload():
    import gen_inputs
    p = gen_inputs.ensure_synth(W, H, SEED, QUALITY)
    if p is None:
        raise SystemExit("workload file missing and oracle/_ref/stbgen not built: run build() first")
    with open(p, "rb") as f:
        return f.read(), os.path.basename(p)


class ClockSampler:
    """nvidia-smi clock/throttle sampling during the timed region (B200_PROFILING.md)."""
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
                 "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
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
        sm, reasons = [], set()
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1]))
                    out["sm_max_mhz"] = float(f[2])
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
            sm.sort()
            out["sm_mhz"] = sm[len(sm) // 2]
            out["samples"] = len(sm)
        out["reasons"] = sorted(reasons)
        return out


def dist_setup(n_gpus):
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


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of `kernel` per launch from the committed
    `ncu --set full` capture of this workload (profiles/r1_ncu_full_summary.json), else None."""
    try:
        rows = json.load(open(os.path.join(ROOT, "profiles", "r1_ncu_full_summary.json")))
        for r in rows:
            if r["kernel"] == kernel:
                m = r["metrics"]
                return int((float(m["dram__bytes_read.sum"]) + float(m["dram__bytes_write.sum"])) * 1e6)
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
    return {"value": round(n * MEGAPIXELS / dt, 2), "unit": "MP/s", "cores": 1, "kind": "port",
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

            n = steps if backend == "gpu_hybrid" else max(5, min(steps, 10))
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


def run_ours(args):
    import torch
    from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes
    rank, world, local = dist_setup(args.gpus)
    data, fname = load_workload()
    sess = DecodeSession(f"cuda:{local}")
    host = pinned_bytes(data)
    info = sess.parse(host)
    assert (info.sizes_x[0], info.sizes_y[0]) == (W, H)
    planes = sess.alloc_planes()
    st = sess.dec.scan_stats(0)
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=sess.device)  # > 126 MB L2
    sample = torch.empty(4096, dtype=torch.uint8, pin_memory=True)
    stream = torch.cuda.current_stream()

    def flush_l2():
        flush.fill_(1)

    # ---- device-resident throughput ("value") ---------------------------------------------
    sess.transfer()
    for _ in range(args.warmup):
        sess.decode(planes)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier(world)
    for a, b in ev:
        flush_l2()
        a.record(stream)
        sess.decode(planes)
        b.record(stream)
    barrier(world)
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    dev_ms = max_over_ranks(dev_ms, world)

    # ---- end to end through the C ABI with host input ("e2e") -----------------------------
    def e2e_step():
        sess.parse(host)
        sess.transfer()
        sess.decode(planes)
        sample.copy_(planes[0].view(-1)[:4096], non_blocking=True)
        torch.cuda.synchronize()

    barrier(world)
    e2e_times = e2e_loop(e2e_step, flush_l2, args.steps, args.warmup)
    barrier(world)
    clocks = sampler.stop() if rank == 0 else None
    e2e_s = max_over_ranks(sum(e2e_times), world)

    # ---- pipelined throughput: several images in flight on one GPU (extra, not the headline) --
    # `depth` decoders, each with its own scratch, planes and stream, decode the same resident
    # image round-robin; kernels of different images overlap on the SMs that a single image's
    # latency-bound Huffman stage leaves idle.  Inputs resident (transfer done), device timed.
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

    # ---- per-kernel times (profiling pass, separate from the timed regions) ---------------
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

    # ---- optional RGB output (extra, not the headline): the same image decoded
    # straight to interleaved RGB by the fused IDCT kernel, device timed, L2 flushed -----------
    rgb = torch.empty((H, W, 3), dtype=torch.uint8, device=sess.device)
    for _ in range(3):
        sess.decode_rgbi(rgb)
    rgb_ms = 0.0
    for _ in range(reps):
        flush_l2()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        sess.decode_rgbi(rgb)
        b.record(stream)
        torch.cuda.synchronize()
        rgb_ms += a.elapsed_time(b)
    rgb_ms /= reps

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
    own = rows
    top = max(own, key=lambda k: own[k]["ms"])
    achieved = own[top]["gbs"] or 0.0
    total_kernel_ms = sum(v["ms"] for v in rows.values())
    ms_per_step = dev_ms / args.steps
    value = world * MEGAPIXELS / (ms_per_step * 1e-3)
    e2e_value = world * args.steps * MEGAPIXELS / e2e_s
    cpu = cpu_baseline_port(data)
    rp = reference_parse_us(data)
    if rp is not None:
        cpu["reference_parse_us"] = rp
        cpu["reference_parse_note"] = "reference reader.cpp parse_header of the workload, 1 thread"
    out = {
        "metric": "decode throughput, 12MP 4:2:0 baseline JPEG", "value": round(value, 1),
        "unit": "MP/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(ms_per_step, 5), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int16/int32 fixed point", "data": "synthetic",
        "images_per_s": round(value / MEGAPIXELS, 1),
        "config": {"workload": f"{fname}: {W}x{H} 4:2:0 q90 stb_image_write, {len(data)} B, "
                               f"{st.num_segments} segment(s), {st.num_subsequences} subsequences "
                               f"of {st.subsequence_bytes} B",
                   "l2": "flushed (512 MiB fill) before every timed step",
                   "parallelism": f"replicas x{world}"},
        "e2e": {"value": round(e2e_value, 1), "unit": "MP/s",
                "ms_per_image": round(e2e_s / args.steps * 1e3, 5),
                "h2d_bytes_per_step": sess.dec.transfer_size(), "d2h_bytes_per_step": 4096,
                "protocol": "parse_header+get_buffer_size+transfer+decode+4KiB D2H+sync, wall clock"},
        "throughput_pipelined": {"value": round(pipe_value, 1), "unit": "MP/s", "images_in_flight": depth,
                                 "images": rounds * depth * world,
                                 "note": "same image, 4 decoders/streams per GPU, inputs resident, "
                                         "CUDA events; L2 not flushed between images"},
        "rgb_output": {"value": round(MEGAPIXELS / (rgb_ms * 1e-3), 1), "unit": "MP/s",
                       "ms_per_image": round(rgb_ms, 5), "steps": reps,
                       "note": "jpeggpu_b200_decode_rgbi: same scan kernels, IDCT fused with chroma "
                               "replication + YCbCr->RGB (bit-exact vs the reference's conv_to_rgbi); "
                               "inputs resident, CUDA events, L2 flushed; one GPU"},
        "gpu_launches": launches * args.steps,
        "launches_per_image": launches,
        "roofline": {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": round(achieved / peak, 5), "traffic": ncu_traffic(top),
                     "peak_source": peak_src,
                     "traffic_source": "profiles/r1_ncu_full_summary.json (ncu --set full, cold caches; the "
                                       "zero-fill stays in L2)",
                     "note": "Huffman kernels are issue/latency bound (bit-serial); see stages"},
        "stages": rows,
        "sum_kernel_ms": round(total_kernel_ms, 5),
        "cpu_baseline": cpu,
        "clocks": clocks,
    }
    emit(out)


def run_reference(args):
    """The reference arm: the UNMODIFIED reference decoder (oracle/_ref/libjpeggpu_ref.so) driven
    through its own public API with the e2e protocol; CPU oracle port if it cannot run."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    data, fname = load_workload()
    base = {"impl": "reference", "metric": "decode throughput, 12MP 4:2:0 baseline JPEG",
            "unit": "MP/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "data": "synthetic",
            "config": {"workload": f"{fname}: {W}x{H} 4:2:0 q90 stb_image_write, {len(data)} B"}}
    gpu_ok = False
    try:
        import torch
        import ref_bindings as rb
        gpu_ok = torch.cuda.is_available() and rb.available()
    except Exception:
        gpu_ok = False
    if not gpu_ok:
        cpu = cpu_baseline_port(data, budget_s=20.0)
        base.update({"value": cpu["value"], "ms_per_step": round(MEGAPIXELS / cpu["value"] * 1e3, 3),
                     "cpu_baseline": cpu, "dtype": "int16/int32 fixed point",
                     "e2e": {"value": cpu["value"], "unit": "MP/s", "h2d_bytes_per_step": 0,
                             "d2h_bytes_per_step": 0},
                     "note": "reference CUDA library unavailable: timed the CPU oracle port"})
        emit(base)
        return
    import numpy as np
    import torch
    import ref_bindings as rb
    torch.cuda.set_device(0)
    host = torch.empty(len(data), dtype=torch.uint8, pin_memory=True)
    host.numpy()[:] = np.frombuffer(data, dtype=np.uint8)
    rd = rb.RefDecoder()
    st, info = rd.parse(host.data_ptr(), len(data))
    assert st == 0, st
    planes = [torch.zeros((info.sizes_y[c], info.sizes_x[c]), dtype=torch.uint8, device="cuda")
              for c in range(info.num_components)]
    need = rd.buffer_size()
    tmp = torch.empty(need, dtype=torch.uint8, device="cuda")
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    sample = torch.empty(4096, dtype=torch.uint8, pin_memory=True)
    s = torch.cuda.current_stream().cuda_stream
    ptrs = [p.data_ptr() for p in planes]
    pitches = [p.stride(0) for p in planes]

    def step():
        st, _ = rd.parse(host.data_ptr(), len(data))
        n = rd.buffer_size()
        assert st == 0 and n <= need
        assert rd.transfer(tmp.data_ptr(), need, s) == 0
        assert rd.decode(ptrs, pitches, tmp.data_ptr(), need, s) == 0
        sample.copy_(planes[0].view(-1)[:4096], non_blocking=True)
        torch.cuda.synchronize()

    sampler = ClockSampler(0)
    sampler.start()
    times = e2e_loop(step, lambda: flush.fill_(1), args.steps, args.warmup)
    # device-only time of the reference's decode call (transfer done), CUDA events
    ev = []
    for _ in range(args.steps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        assert rd.decode(ptrs, pitches, tmp.data_ptr(), need, s) == 0
        b.record()
        ev.append((a, b))
    torch.cuda.synchronize()
    clocks = sampler.stop()
    dev_ms = sum(a.elapsed_time(b) for a, b in ev) / args.steps
    rd.close()
    nvj = nvjpeg_leg(host, len(data), lambda: flush.fill_(1), args.steps, args.warmup)
    total = sum(times)
    value = args.steps * MEGAPIXELS / total
    cpu = {"value": round(value, 1), "unit": "MP/s", "cores": 1, "kind": "reference",
           "sample": "the reference has no CPU decode path: this line times its CUDA decoder "
                     "(oracle/_ref/libjpeggpu_ref.so, built from /root/reference/src for sm_100) "
                     "driven by 1 host thread; host-side parse is inside the timed region"}
    rp = reference_parse_us(data)
    if rp is not None:
        cpu["reference_parse_us"] = rp
    base.update({"value": round(value, 1), "ms_per_step": round(total / args.steps * 1e3, 5),
                 "images_per_s": round(args.steps / total, 1),
                 "dtype": "int16/int32 fixed point",
                 "device_only": {"value": round(MEGAPIXELS / (dev_ms * 1e-3), 1), "unit": "MP/s",
                                 "ms_per_step": round(dev_ms, 5),
                                 "note": "CUDA events around jpeggpu_decoder_decode only, L2 flushed"},
                 "e2e": {"value": round(value, 1), "unit": "MP/s", "h2d_bytes_per_step": len(data) + 4 * 64 + 4 * 7168,
                         "d2h_bytes_per_step": 4096,
                         "protocol": "parse_header+get_buffer_size+transfer+decode+4KiB D2H+sync, wall clock"},
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
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
