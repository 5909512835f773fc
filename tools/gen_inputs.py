#!/usr/bin/env python
"""Generate / locate the input files BASELINE.json names.

Files live in data/_gen/ (git-ignored, travels to the GPU box with the gpurun snapshot):
  IMG_6510.JPG          config 1: copied from /root/reference/images when present
  synth_<W>x<H>_s<seed>_q<q>.jpg   stb_image_write output of the deterministic noise picture,
                        written by oracle/_ref/stbgen (built from the reference's bundled
                        util/stb_image_write.h by oracle/Makefile).
Nothing here is on the product path; it only prepares bytes to feed to the decoder.
"""
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GEN = os.path.join(ROOT, "data", "_gen")
STBGEN = os.path.join(ROOT, "oracle", "_ref", "stbgen")
REF_IMG = "/root/reference/images/IMG_6510.JPG"


def synth_path(w, h, seed, quality=90, amp=1.0):
    tag = f"synth_{w}x{h}_s{seed}_q{quality}" + ("" if amp == 1.0 else f"_a{amp:g}")
    return os.path.join(GEN, tag + ".jpg")


def ensure_synth(w, h, seed, quality=90, amp=1.0):
    """Returns the path of the stb-encoded synthetic picture, generating it if needed."""
    p = synth_path(w, h, seed, quality, amp)
    if not os.path.exists(p):
        if not os.path.exists(STBGEN):
            return None
        os.makedirs(GEN, exist_ok=True)
        tmp = p + f".tmp{os.getpid()}"
        subprocess.run([STBGEN, tmp, str(w), str(h), str(seed), str(quality), f"{amp:g}"], check=True)
        os.replace(tmp, p)
    return p


def ensure_img6510():
    p = os.path.join(GEN, "IMG_6510.JPG")
    if not os.path.exists(p):
        if not os.path.exists(REF_IMG):
            return None
        os.makedirs(GEN, exist_ok=True)
        shutil.copyfile(REF_IMG, p)
    return p


def main():
    made = [ensure_img6510(),
            ensure_synth(4032, 3024, 6510),      # config 2
            ensure_synth(3008, 2000, 1),         # config 3 (4:2:0)
            ensure_synth(3008, 2000, 2, 95),     # config 3 (4:4:4)
            ensure_synth(3001, 1999, 3),         # non-aligned twin
            ensure_synth(1920, 1080, 1),         # config 5 sample
            ensure_synth(8064, 6048, 48)]        # config 4
    for m in made:
        print(m, os.path.getsize(m) if m else "unavailable")
    return 0


if __name__ == "__main__":
    sys.exit(main())
