"""ctypes bindings of oracle/_ref/libnvjpeg_harness.so (oracle/nvjpeg_harness.c): nvJPEG from the
local CUDA toolkit, decoupled API, NVJPEG_OUTPUT_UNCHANGED.  TEST / BASELINE INFRASTRUCTURE: a
third, independent decoder for pixel cross-checks (max-abs <= 2, different IDCT) and the second
baseline BASELINE.json names.  Never on the product path."""
import ctypes as C
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libnvjpeg_harness.so")

BACKENDS = {"default": 0, "hybrid": 1, "gpu_hybrid": 2, "hardware": 3}


def available():
    return os.path.exists(LIB)


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB)
        L.nvjh_create.restype = C.c_void_p
        L.nvjh_create.argtypes = [C.c_int, C.POINTER(C.c_int)]
        L.nvjh_destroy.argtypes = [C.c_void_p]
        L.nvjh_info.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_int),
                                C.POINTER(C.c_int * 4), C.POINTER(C.c_int * 4)]
        L.nvjh_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p * 4),
                                  C.POINTER(C.c_int * 4), C.c_void_p, C.c_int]
        _lib = L
    return _lib


class NvJpeg:
    """One nvJPEG decoder of the given backend; raises RuntimeError if nvJPEG refuses it."""

    def __init__(self, backend="gpu_hybrid"):
        self.L = lib()
        st = C.c_int(0)
        self.h = self.L.nvjh_create(BACKENDS[backend], C.byref(st))
        if not self.h:
            raise RuntimeError(f"nvJPEG backend {backend} unavailable (status {st.value})")
        self.backend = backend

    def info(self, ptr, size):
        ch = C.c_int()
        w = (C.c_int * 4)()
        h = (C.c_int * 4)()
        st = self.L.nvjh_info(self.h, ptr, size, C.byref(ch), C.byref(w), C.byref(h))
        if st:
            raise RuntimeError(f"nvjpegGetImageInfo status {st}")
        return ch.value, list(w), list(h)

    def decode(self, ptr, size, plane_ptrs, pitches, stream=0, sync=True):
        pp = (C.c_void_p * 4)(*(list(plane_ptrs) + [None] * (4 - len(plane_ptrs))))
        pt = (C.c_int * 4)(*(list(pitches) + [0] * (4 - len(pitches))))
        return self.L.nvjh_decode(self.h, ptr, size, C.byref(pp), C.byref(pt), stream, int(sync))

    def close(self):
        if self.h:
            self.L.nvjh_destroy(self.h)
            self.h = None
