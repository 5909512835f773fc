#!/usr/bin/env python
"""Print the interesting numbers of a pair of bench lines: show_bench.py <ours.json> [<reference.json>]"""
import json
import sys


def load(p):
    for line in open(p):
        line = line.strip()
        if line.startswith("{"):
            return json.loads(line)
    return None


def main():
    o = load(sys.argv[1])
    r = load(sys.argv[2]) if len(sys.argv) > 2 else None
    def row(name, a, b):
        ratio = f"{a / b:.2f}x" if a and b else "-"
        print(f"{name:48s} ours {a!s:>12}  ref {b!s:>12}  {ratio}")
    row("value MP/s (device resident)", o["value"], r and r.get("value"))
    row("e2e MP/s", o["e2e"]["value"], r and r["e2e"]["value"])
    print(f"{'ms_per_step / e2e ms':48s} ours {o['ms_per_step']} / {o['e2e']['ms_per_image']}" +
          (f"   ref {r['ms_per_step']} / {r['e2e']['ms_per_image']}" if r else ""))
    row("sustained MP/s", o["sustained"]["value"], r and r.get("sustained", {}).get("value"))
    for k, v in (o.get("configs") or {}).items():
        rv = (r or {}).get("configs", {}).get(k, {})
        print(f"  {k:46s} ours dev {v['device_ms']:.4f} e2e {v['e2e_ms']:.4f} ms" +
              (f"   ref dev {rv['device_ms']:.4f} e2e {rv['e2e_ms']:.4f} ms   x{rv['device_ms'] / v['device_ms']:.2f} / x{rv['e2e_ms'] / v['e2e_ms']:.2f}" if "device_ms" in rv else ""))
    b, rb = o.get("batch"), (r or {}).get("batch")
    if b:
        print("batch:", {k: b[k] for k in b if k in ("images_per_s", "seconds", "images", "host_busy_seconds_slowest_thread", "host_threads_per_gpu", "path", "matches_per_image_path", "unavailable")})
    if rb:
        print("ref batch:", {k: rb[k] for k in rb if k in ("images_per_s", "seconds", "images")})
    print("stages:", {k: v["ms"] for k, v in o["stages"].items()})
    print("roofline:", {k: o["roofline"][k] for k in ("kernel", "achieved", "frac")})
    print("clocks:", o.get("clocks"))


if __name__ == "__main__":
    main()
