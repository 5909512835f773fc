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


# Named inputs of the golden fixtures (tests/golden/reference_digests.json, tools/make_golden.py).
def four_scan_6mp():
    """BASELINE config 3 at its stated size: 3008x2000, four components, one scan each, DHT before
    every SOS, 1x1 sampling (the one non-interleaved structure the reference decodes).
    Written by tests/jpeg_transcode.py from the coefficient planes of Pillow's interleaved CMYK
    file; the transcode takes ~20 s of Python, so the file is kept under data/_gen."""
    import jpeg_transcode as jt
    import oracle_bindings as ob
    path = os.path.join(gen_inputs.GEN, "pil_cmyk_3008x2000_s39_four_scans.jpg")
    if os.path.exists(path):
        return read(path)
    src = pillow_jpeg(w=3008, h=2000, seed=39, mode="CMYK")
    ref = ob.decode(src)
    assert ref["status"] == 0
    data = jt.transcode(src, ref["planes"], [[0], [1], [2], [3]], restart_interval=0, dht_every_scan=True)
    os.makedirs(gen_inputs.GEN, exist_ok=True)
    tmp = path + f".tmp{os.getpid()}"
    with open(tmp, "wb") as f:
        f.write(data)
    os.replace(tmp, path)
    return data


NAMED = {
    "IMG_6510": lambda: img6510(),
    "stb_4032x3024_s6510_q90": lambda: synth(4032, 3024, 6510, 90),
    "stb_3008x2000_s1_q90": lambda: synth(3008, 2000, 1, 90),
    "stb_3008x2000_s2_q95": lambda: synth(3008, 2000, 2, 95),
    "stb_3001x1999_s3_q90": lambda: synth(3001, 1999, 3, 90),
    "stb_1920x1080_s1_q90": lambda: synth(1920, 1080, 1, 90),
    # BASELINE config 4 (48 MP, one restart segment) and its harder twin: amplitudes x8 make the
    # scan 18.3 MB, more than one sync_subsequences CTA of the reference covers
    "pil_cmyk_3008x2000_four_scans": four_scan_6mp,
    "stb_8064x6048_s48_q90": lambda: synth(8064, 6048, 48, 90),
    "stb_8064x6048_s48_q90_a8": lambda: synth(8064, 6048, 48, 90, 8.0),
    "pil_422_3008x2000": lambda: pillow_jpeg(w=3008, h=2000, seed=12, mode="YCbCr", subsampling=1),
    "pil_444_1504x1000": lambda: pillow_jpeg(w=1504, h=1000, seed=11, mode="YCbCr", subsampling=0),
    "pil_gray_3008x2000": lambda: pillow_jpeg(w=3008, h=2000, seed=14, mode="L"),
    "pil_420_rst63_1008x752": lambda: pillow_jpeg(w=1008, h=752, seed=17, mode="YCbCr",
                                                   subsampling=2, restart_blocks=63),
    "pil_420_rst1_333x211": lambda: pillow_jpeg(w=333, h=211, seed=18, mode="YCbCr",
                                                 subsampling=2, restart_blocks=1),
    "pil_420_opt_1280x720": lambda: pillow_jpeg(w=1280, h=720, seed=19, mode="YCbCr",
                                                 subsampling=2, optimize=True),
}


def read_named(name):
    fn = NAMED.get(name)
    return fn() if fn else None
