"""Builds and binds tests/host_emul/emul.cpp: the product's host+device shared code (parser, LUT
builder, Huffman thread bodies) run sequentially on the CPU.  TEST INFRASTRUCTURE."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "host_emul", "emul.cpp")
CSRC = os.path.join(ROOT, "jpeggpu_b200", "csrc")
DEFS = os.environ.get("JB2_DEFS", "").split()  # e.g. "-DJB2_SUB_BYTES=32 -DJB2_OVERLAP_SUBS=12" (experiments)
OUT = os.path.join(ROOT, "tests", "host_emul", "_build",
                   "libemul" + "".join(d.replace("-D", "_").replace("=", "") for d in DEFS) + ".so")
CUDA_INC = "/usr/local/cuda/include"

_lib = None


def lib():
    global _lib
    if _lib is None:
        deps = [SRC] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".hpp", ".cpp"))]
        if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in deps):
            os.makedirs(os.path.dirname(OUT), exist_ok=True)
            subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", *DEFS, "-I" + CSRC,
                            "-I" + os.path.join(ROOT, "include"), "-I" + CUDA_INC, SRC,
                            os.path.join(CSRC, "parser.cpp"), "-lpthread", "-o", OUT], check=True)
        L = C.CDLL(OUT)
        L.emul_parse.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        L.emul_decode_scan.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_size_t,
                                       C.c_void_p, C.c_size_t, C.c_void_p]
        L.emul_pair_equivalence.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_long)]
        L.emul_walk_compare.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_int),
                                        C.POINTER(C.c_int)]
        _lib = L
    return _lib


def walk_compare(data: bytes, threads: bool):
    """(difference code, pool use, status): eager walk vs deferred walk (+ helper threads).
    pool use: 0 not kicked, 1 kicked but the result was discarded (sequential re-walk), 2 kept."""
    buf = np.frombuffer(data, dtype=np.uint8)
    used, st = C.c_int(), C.c_int()
    rc = lib().emul_walk_compare(buf.ctypes.data, len(data), int(threads), C.byref(used), C.byref(st))
    return rc, used.value, st.value


def pair_equivalence(data: bytes, tries=3):
    """(records that differ between pair stepping and single stepping, pair entries in the LUTs)."""
    buf = np.frombuffer(data, dtype=np.uint8)
    pairs = C.c_long()
    bad = lib().emul_pair_equivalence(buf.ctypes.data, len(data), tries, C.byref(pairs))
    return bad, pairs.value


def parse(data: bytes):
    buf = np.frombuffer(data, dtype=np.uint8)
    out = np.zeros(16, dtype=np.int32)
    st = lib().emul_parse(buf.ctypes.data, len(data), out.ctypes.data)
    return st, out


def decode_scan(data: bytes, scan, coef_count, destuffed_bytes=0, corrupt_records=False):
    buf = np.frombuffer(data, dtype=np.uint8)
    coef = np.zeros(coef_count, dtype=np.int16)
    stats = np.zeros(8, dtype=np.int32)
    if corrupt_records:
        stats[7] = 12345  # emul.cpp test hook: damage sync records before huff_link
    dst = np.zeros(destuffed_bytes, dtype=np.uint8) if destuffed_bytes else None
    st = lib().emul_decode_scan(buf.ctypes.data, len(data), scan, coef.ctypes.data, coef.size,
                                dst.ctypes.data if dst is not None else None, destuffed_bytes,
                                stats.ctypes.data)
    return st, coef, stats, dst
