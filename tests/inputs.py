"""Input files for the tests: the in-repo photo, stb synthetics (data/_gen, see
tools/gen_inputs.py) and Pillow-encoded variants generated on the fly (libjpeg-turbo encoder:
other sampling factors, restart intervals, grayscale, optimised tables)."""
import io
import os

import numpy as np

import gen_inputs


def read(path):
    with open(path, "rb") as f:
        return f.read()


def img6510():
    p = gen_inputs.ensure_img6510()
    return read(p) if p else None


def synth(w, h, seed, quality=90, amp=1.0):
    p = gen_inputs.ensure_synth(w, h, seed, quality, amp)
    return read(p) if p else None


def noise_rgb(w, h, seed):
    """Smooth-ish deterministic RGB picture (numpy only)."""
    rng = np.random.RandomState(seed)
    img = np.zeros((h, w, 3), dtype=np.float32)
    for cell, amp in ((64, 90.0), (16, 50.0), (4, 25.0), (1, 10.0)):
        gh, gw = h // cell + 2, w // cell + 2
        g = rng.rand(gh, gw, 3).astype(np.float32) - 0.5
        up = np.kron(g, np.ones((cell, cell, 1), dtype=np.float32))[:h, :w]
        img += amp * up
    return np.clip(img + 128.0, 0, 255).astype(np.uint8)


def pillow_jpeg(w, h, seed, mode="YCbCr", subsampling=2, quality=85, restart_blocks=0,
                optimize=False):
    """mode: 'YCbCr' (3 comp), 'L' (gray), 'CMYK' (4 comp).  subsampling: 0=4:4:4, 1=4:2:2,
    2=4:2:0 (Pillow's numbering)."""
    from PIL import Image
    rgb = noise_rgb(w, h, seed)
    if mode == "L":
        im = Image.fromarray(rgb[:, :, 0], "L")
    elif mode == "CMYK":
        four = np.concatenate([rgb, rgb[:, :, :1][:, ::-1]], axis=2)
        im = Image.fromarray(four, "CMYK")
    else:
        im = Image.fromarray(rgb, "RGB")
    buf = io.BytesIO()
    kw = dict(quality=quality, optimize=optimize)
    if mode == "YCbCr":
        kw["subsampling"] = subsampling
    if restart_blocks:
        kw["restart_marker_blocks"] = restart_blocks
    im.save(buf, "JPEG", **kw)
    return buf.getvalue()
