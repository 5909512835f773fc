"""ctypes bindings of the CPU oracle (oracle/jpeg_oracle.c).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB_PATH = os.path.join(ORACLE_DIR, "_build", "libjpeg_oracle.so")


class ScanInfo(C.Structure):
    _fields_ = [
        ("num_components", C.c_int),
        ("comp_idx", C.c_int * 4),
        ("dc_tbl", C.c_int * 4),
        ("ac_tbl", C.c_int * 4),
        ("begin", C.c_int),
        ("end", C.c_int),
        ("dus_per_mcu", C.c_int),
        ("mcus_x", C.c_int),
        ("mcus_y", C.c_int),
        ("data_size_x", C.c_int * 4),
        ("data_size_y", C.c_int * 4),
        ("num_segments", C.c_int),
        ("restart_interval", C.c_int),
    ]


class Info(C.Structure):
    _fields_ = [
        ("width", C.c_int),
        ("height", C.c_int),
        ("num_components", C.c_int),
        ("comp_id", C.c_int * 4),
        ("ssx", C.c_int * 4),
        ("ssy", C.c_int * 4),
        ("qidx", C.c_int * 4),
        ("size_x", C.c_int * 4),
        ("size_y", C.c_int * 4),
        ("ss_max_x", C.c_int),
        ("ss_max_y", C.c_int),
        ("restart_interval", C.c_int),
        ("num_scans", C.c_int),
        ("scans", ScanInfo * 4),
        ("qtables", (C.c_uint8 * 64) * 4),
    ]


_lib = None


def build():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "_build/libjpeg_oracle.so"], check=True)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < os.path.getmtime(
            os.path.join(ORACLE_DIR, "jpeg_oracle.c")
        ):
            build()
        L = C.CDLL(LIB_PATH)
        L.orc_parse.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Info)]
        L.orc_decode.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Info), C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_int]
        L.orc_stream_coef_count.restype = C.c_size_t
        L.orc_stream_coef_count.argtypes = [C.POINTER(Info), C.c_int]
        L.orc_plane_coef_count.restype = C.c_size_t
        L.orc_plane_coef_count.argtypes = [C.POINTER(Info), C.c_int, C.POINTER(C.c_int)]
        L.orc_destuff_bound.restype = C.c_size_t
        L.orc_destuff_bound.argtypes = [C.c_size_t, C.c_int]
        L.orc_destuff.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p,
                                  C.POINTER(C.c_size_t), C.c_void_p, C.c_void_p, C.c_int]
        L.orc_idct_block.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.orc_idct_block.restype = None
        L.orc_count_symbols.restype = C.c_long
        L.orc_count_symbols.argtypes = [C.c_void_p, C.c_size_t]
        L.orc_conv_to_rgbi.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def parse(data: bytes):
    info = Info()
    buf = np.frombuffer(data, dtype=np.uint8)
    st = lib().orc_parse(buf.ctypes.data, len(data), C.byref(info))
    return st, info


def decode(data: bytes, ref_signed_q=False):
    """Returns dict(status, info, stream=[per scan int16], planes=[per comp int16 2D],
    pixels=[per comp uint8 2D])."""
    st, info = parse(data)
    if st != 0:
        return {"status": st, "info": info}
    L = lib()
    buf = np.frombuffer(data, dtype=np.uint8)
    stream = [np.zeros(L.orc_stream_coef_count(C.byref(info), s), dtype=np.int16)
              for s in range(info.num_scans)]
    planes, pixels = [], []
    for c in range(info.num_components):
        stride = C.c_int()
        cnt = L.orc_plane_coef_count(C.byref(info), c, C.byref(stride))
        planes.append(np.zeros((cnt // stride.value, stride.value), dtype=np.int16))
        pixels.append(np.zeros((info.size_y[c], info.size_x[c]), dtype=np.uint8))
    sp = (C.c_void_p * 4)(*[a.ctypes.data for a in stream] + [None] * (4 - len(stream)))
    pp = (C.c_void_p * 4)(*[a.ctypes.data for a in planes] + [None] * (4 - len(planes)))
    xp = (C.c_void_p * 4)(*[a.ctypes.data for a in pixels] + [None] * (4 - len(pixels)))
    pitch = (C.c_int * 4)(*[a.shape[1] for a in pixels] + [0] * (4 - len(pixels)))
    st = L.orc_decode(buf.ctypes.data, len(data), C.byref(info), sp, pp, xp, pitch,
                      1 if ref_signed_q else 0)
    return {"status": st, "info": info, "stream": stream, "planes": planes, "pixels": pixels}


def destuff(scan_bytes: bytes, subseq_bytes: int):
    L = lib()
    n = len(scan_bytes)
    src = np.frombuffer(scan_bytes, dtype=np.uint8)
    out = np.zeros(L.orc_destuff_bound(n, subseq_bytes), dtype=np.uint8)
    max_segs = n // 2 + 2
    seg_off = np.zeros(max_segs, dtype=np.uint64)
    seg_len = np.zeros(max_segs, dtype=np.uint64)
    used = C.c_size_t()
    k = L.orc_destuff(src.ctypes.data, n, subseq_bytes, out.ctypes.data, C.byref(used),
                      seg_off.ctypes.data, seg_len.ctypes.data, max_segs)
    assert k >= 0
    return out[: used.value], seg_off[:k].astype(np.int64), seg_len[:k].astype(np.int64)


def idct_block(coef, q, ref_signed_q=False):
    coef = np.ascontiguousarray(coef, dtype=np.int16)
    q = np.ascontiguousarray(q, dtype=np.uint8)
    out = np.zeros((8, 8), dtype=np.uint8)
    lib().orc_idct_block(coef.ctypes.data, q.ctypes.data, out.ctypes.data, 8, int(ref_signed_q))
    return out


REF_UTIL_LIB = os.path.join(ORACLE_DIR, "_ref", "libref_util.so")


def _conv(fn, planes, ssx, ssy):
    """planes: 1 or 3 dense uint8 2-D arrays.  Returns (rc, HxWx3 uint8)."""
    planes = [np.ascontiguousarray(p, dtype=np.uint8) for p in planes]
    n = len(planes)
    sx = (C.c_int * 4)(*[p.shape[1] for p in planes] + [0] * (4 - n))
    sy = (C.c_int * 4)(*[p.shape[0] for p in planes] + [0] * (4 - n))
    fx = (C.c_int * 4)(*(list(ssx) + [0] * (4 - len(ssx))))
    fy = (C.c_int * 4)(*(list(ssy) + [0] * (4 - len(ssy))))
    out = np.zeros((planes[0].shape[0], planes[0].shape[1], 3), dtype=np.uint8)
    ptr = [p.ctypes.data for p in planes] + [None] * (3 - n)
    rc = fn(sx, sy, n, fx, fy, ptr[0], ptr[1], ptr[2], out.ctypes.data)
    return rc, out


def conv_to_rgbi(planes, ssx, ssy):
    """orc_conv_to_rgbi (restatement of the reference's util/util.h conv_to_rgbi)."""
    return _conv(lib().orc_conv_to_rgbi, planes, ssx, ssy)


def ref_util_available():
    return os.path.exists(REF_UTIL_LIB)


def ref_conv_to_rgbi(planes, ssx, ssy):
    """The reference's own conv_to_rgbi, compiled from its header by oracle/Makefile."""
    L = C.CDLL(REF_UTIL_LIB)
    L.ref_conv_to_rgbi.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    return _conv(L.ref_conv_to_rgbi, planes, ssx, ssy)
