"""jpeggpu_b200 — Python host-side mirror of the jpeggpu C API over libjpeggpu_b200.so.

The product is the C-ABI shared library (include/jpeggpu/jpeggpu.h, include/jpeggpu_b200.h) built
from jpeggpu_b200/csrc for sm_100a.  This module only binds it with ctypes so tests and bench.py
can drive it the way reference callers do (example/example_tool.c:101-128):
startup -> parse_header -> get_buffer_size -> transfer -> decode -> (caller syncs) -> cleanup.
There is no CPU fallback: if the library is missing, importing the bindings fails loudly.
"""
import ctypes as C
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "jpeggpu_b200", "csrc")
LIB_PATH = os.environ.get("JPEGGPU_B200_LIB") or os.path.join(ROOT, "jpeggpu_b200", "lib",
                                                             "libjpeggpu_b200.so")

MAX_COMP = 4

(SUCCESS, INVALID_ARGUMENT, INVALID_JPEG, INTERNAL_ERROR, NOT_SUPPORTED, OUT_OF_HOST_MEMORY,
 INCOMPLETE_BITSTREAM) = range(7)

STATUS_NAMES = ["JPEGGPU_SUCCESS", "JPEGGPU_INVALID_ARGUMENT", "JPEGGPU_INVALID_JPEG",
                "JPEGGPU_INTERNAL_ERROR", "JPEGGPU_NOT_SUPPORTED", "JPEGGPU_OUT_OF_HOST_MEMORY",
                "JPEGGPU_INCOMPLETE_BITSTREAM"]

# every symbol include/jpeggpu/jpeggpu.h and include/jpeggpu_b200.h declare
API_SYMBOLS = [
    "jpeggpu_get_status_string", "jpeggpu_decoder_startup", "jpeggpu_set_logging", "is_css_444",
    "jpeggpu_decoder_parse_header", "jpeggpu_decoder_get_buffer_size", "jpeggpu_decoder_transfer",
    "jpeggpu_decoder_decode", "jpeggpu_decoder_cleanup",
    "jpeggpu_b200_get_coefficient_size", "jpeggpu_b200_decode_coefficients",
    "jpeggpu_b200_get_scan_stats", "jpeggpu_b200_get_transfer_size",
    "jpeggpu_b200_set_profiling", "jpeggpu_b200_get_kernel_times", "jpeggpu_b200_kernel_name",
    "jpeggpu_b200_decode_batch", "jpeggpu_b200_parse_batch", "jpeggpu_b200_set_deferred_walk",
    "jpeggpu_b200_decode_rgbi", "jpeggpu_b200_convert_rgbi", "jpeggpu_b200_read_scan_info",
    "jpeggpu_b200_batch_create", "jpeggpu_b200_batch_destroy", "jpeggpu_b200_batch_plan",
    "jpeggpu_b200_batch_decode", "jpeggpu_b200_batch_read_scan_info",
    "jpeggpu_b200_batch_set_host_arena", "jpeggpu_b200_destuff_static",
]

MAX_KERNELS = 16


class Subsampling(C.Structure):
    _fields_ = [("x", C.c_int * MAX_COMP), ("y", C.c_int * MAX_COMP)]


class ImgInfo(C.Structure):
    _fields_ = [("sizes_x", C.c_int * MAX_COMP), ("sizes_y", C.c_int * MAX_COMP),
                ("num_components", C.c_int), ("subsampling", Subsampling)]


class Img(C.Structure):
    _fields_ = [("image", C.c_void_p * MAX_COMP), ("pitch", C.c_int * MAX_COMP)]


class ScanStats(C.Structure):
    _fields_ = [("num_components", C.c_int), ("num_segments", C.c_int), ("dus_per_mcu", C.c_int),
                ("total_mcus", C.c_int), ("scan_bytes", C.c_uint), ("destuffed_bytes", C.c_uint),
                ("num_subsequences", C.c_uint), ("num_data_units", C.c_uint),
                ("subsequence_bytes", C.c_int), ("kernel_launches", C.c_int),
                ("off_control", C.c_size_t), ("off_destuffed", C.c_size_t),
                ("off_coefficients", C.c_size_t), ("off_sync_records", C.c_size_t),
                ("off_slot_prefix", C.c_size_t), ("off_segments", C.c_size_t),
                ("off_scan_info", C.c_size_t)]


class ScanInfo(C.Structure):
    """struct jpeggpu_b200_scan_info (include/jpeggpu_b200.h)"""
    _fields_ = [("num_subsequences", C.c_uint), ("num_segments", C.c_uint), ("scan_end", C.c_uint),
                ("status", C.c_uint)]


DESTUFF_STATIC_MAX_BYTES = 64 << 20  # JPEGGPU_B200_DESTUFF_STATIC_MAX_BYTES


class BatchItem(C.Structure):
    """struct jpeggpu_b200_batch_item (include/jpeggpu_b200.h)"""
    _fields_ = [("data", C.c_void_p), ("size", C.c_size_t), ("img", Img), ("info", ImgInfo),
                ("status", C.c_int)]


class JpegGpuError(RuntimeError):
    def __init__(self, status, where):
        self.status = status
        name = STATUS_NAMES[status] if 0 <= status < len(STATUS_NAMES) else str(status)
        super().__init__(f"{where}: {name}")


def build_library(defs=""):
    """Compile libjpeggpu_b200.so for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-s", "-C", CSRC, "-j8"]
    if defs:
        cmd.append(f"DEFS={defs}")
    subprocess.run(cmd, check=True)
    return LIB_PATH


_lib = None


def load_library(path=None):
    """dlopen the product library and declare the prototypes. Raises if it is not built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise ImportError(f"{p} is not built: run `python -c 'import __graft_entry__ as g; "
                          "g.build()'` (the decoder has no CPU fallback)")
    L = C.CDLL(p)
    dec = C.c_void_p
    L.jpeggpu_get_status_string.restype = C.c_char_p
    L.jpeggpu_get_status_string.argtypes = [C.c_int]
    L.jpeggpu_decoder_startup.argtypes = [C.POINTER(dec)]
    L.jpeggpu_set_logging.argtypes = [dec, C.c_int]
    L.is_css_444.argtypes = [Subsampling, C.c_int]
    L.jpeggpu_decoder_parse_header.argtypes = [dec, C.POINTER(ImgInfo), C.c_void_p, C.c_size_t]
    L.jpeggpu_decoder_get_buffer_size.argtypes = [dec, C.POINTER(C.c_size_t)]
    L.jpeggpu_decoder_transfer.argtypes = [dec, C.c_void_p, C.c_size_t, C.c_void_p]
    L.jpeggpu_decoder_decode.argtypes = [dec, C.POINTER(Img), C.c_void_p, C.c_size_t, C.c_void_p]
    L.jpeggpu_decoder_cleanup.argtypes = [dec]
    L.jpeggpu_b200_get_coefficient_size.argtypes = [dec, C.POINTER(C.c_int * 4), C.POINTER(C.c_int * 4)]
    L.jpeggpu_b200_decode_coefficients.argtypes = [dec, C.POINTER(C.c_void_p * 4), C.c_void_p,
                                                   C.c_size_t, C.c_void_p]
    L.jpeggpu_b200_decode_rgbi.argtypes = [dec, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p]
    L.jpeggpu_b200_convert_rgbi.argtypes = [C.POINTER(Img), C.POINTER(ImgInfo), C.c_void_p, C.c_int,
                                            C.c_void_p]
    L.jpeggpu_b200_get_scan_stats.argtypes = [dec, C.c_int, C.POINTER(ScanStats), C.POINTER(C.c_int)]
    L.jpeggpu_b200_get_transfer_size.argtypes = [dec, C.POINTER(C.c_size_t)]
    L.jpeggpu_b200_set_profiling.argtypes = [dec, C.c_int]
    L.jpeggpu_b200_set_deferred_walk.argtypes = [dec, C.c_int]
    L.jpeggpu_b200_read_scan_info.argtypes = [dec, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(ScanInfo)]
    L.jpeggpu_b200_destuff_static.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p,
                                              C.c_size_t, C.c_void_p, C.c_void_p]
    L.jpeggpu_b200_batch_create.argtypes = [C.POINTER(C.c_void_p)]
    L.jpeggpu_b200_batch_destroy.argtypes = [C.c_void_p]
    L.jpeggpu_b200_batch_set_host_arena.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    L.jpeggpu_b200_batch_plan.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(C.c_size_t)]
    L.jpeggpu_b200_batch_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t,
                                            C.c_void_p]
    L.jpeggpu_b200_batch_read_scan_info.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                                    C.POINTER(ScanInfo)]
    L.jpeggpu_b200_get_kernel_times.argtypes = [dec, C.POINTER(C.c_float * MAX_KERNELS),
                                                C.POINTER(C.c_int * MAX_KERNELS)]
    L.jpeggpu_b200_kernel_name.restype = C.c_char_p
    L.jpeggpu_b200_kernel_name.argtypes = [C.c_int]
    L.jpeggpu_b200_decode_batch.argtypes = [C.POINTER(BatchItem), C.c_int, C.POINTER(C.c_void_p),
                                            C.POINTER(C.c_void_p), C.POINTER(C.c_size_t), C.c_int]
    L.jpeggpu_b200_parse_batch.argtypes = [C.POINTER(BatchItem), C.c_int, C.c_int,
                                           C.POINTER(C.c_size_t)]
    if path is None:
        _lib = L
    return L


def convert_rgbi(planes, pitches, info, d_rgb, pitch, stream=0, lib=None):
    """jpeggpu_b200_convert_rgbi: device planes (as decode() wrote them) -> interleaved RGB."""
    lib = lib or load_library()
    img = Img()
    for c, (p, pt) in enumerate(zip(planes, pitches)):
        img.image[c] = p
        img.pitch[c] = pt
    _check(lib.jpeggpu_b200_convert_rgbi(C.byref(img), C.byref(info), d_rgb, pitch, stream),
           "jpeggpu_b200_convert_rgbi")


def _check(status, where):
    if status != SUCCESS:
        raise JpegGpuError(status, where)


class Decoder:
    """One jpeggpu_decoder_t.  Methods map 1:1 onto the C entry points; pointers are integers
    (e.g. torch.Tensor.data_ptr(), torch.cuda.Stream.cuda_stream)."""

    def __init__(self, lib=None):
        self.lib = lib or load_library()
        self.handle = C.c_void_p()
        _check(self.lib.jpeggpu_decoder_startup(C.byref(self.handle)), "jpeggpu_decoder_startup")
        self._data_ref = None

    def set_logging(self, on):
        _check(self.lib.jpeggpu_set_logging(self.handle, int(on)), "jpeggpu_set_logging")

    def parse_header(self, data_ptr, size, keepalive=None):
        """data_ptr: host address of the JPEG bytes (borrowed until transfer's copies ran)."""
        info = ImgInfo()
        self._data_ref = keepalive
        _check(self.lib.jpeggpu_decoder_parse_header(self.handle, C.byref(info), data_ptr, size),
               "jpeggpu_decoder_parse_header")
        return info

    def get_buffer_size(self):
        n = C.c_size_t()
        _check(self.lib.jpeggpu_decoder_get_buffer_size(self.handle, C.byref(n)),
               "jpeggpu_decoder_get_buffer_size")
        return n.value

    def transfer(self, d_tmp, tmp_size, stream=0):
        _check(self.lib.jpeggpu_decoder_transfer(self.handle, d_tmp, tmp_size, stream),
               "jpeggpu_decoder_transfer")

    def decode(self, planes, pitches, d_tmp, tmp_size, stream=0):
        img = Img()
        for c, (p, pitch) in enumerate(zip(planes, pitches)):
            img.image[c] = p
            img.pitch[c] = pitch
        _check(self.lib.jpeggpu_decoder_decode(self.handle, C.byref(img), d_tmp, tmp_size, stream),
               "jpeggpu_decoder_decode")

    def coefficient_size(self):
        w, h = (C.c_int * 4)(), (C.c_int * 4)()
        _check(self.lib.jpeggpu_b200_get_coefficient_size(self.handle, C.byref(w), C.byref(h)),
               "jpeggpu_b200_get_coefficient_size")
        return list(w), list(h)

    def decode_coefficients(self, planes, d_tmp, tmp_size, stream=0):
        arr = (C.c_void_p * 4)(*(list(planes) + [None] * (4 - len(planes))))
        _check(self.lib.jpeggpu_b200_decode_coefficients(self.handle, C.byref(arr), d_tmp, tmp_size,
                                                         stream), "jpeggpu_b200_decode_coefficients")

    def decode_rgbi(self, d_rgb, pitch, d_tmp, tmp_size, stream=0):
        """Fused IDCT + chroma replication + YCbCr->RGB (jpeggpu_b200_decode_rgbi)."""
        _check(self.lib.jpeggpu_b200_decode_rgbi(self.handle, d_rgb, pitch, d_tmp, tmp_size, stream),
               "jpeggpu_b200_decode_rgbi")

    def num_scans(self):
        n = C.c_int()
        _check(self.lib.jpeggpu_b200_get_scan_stats(self.handle, 0, None, C.byref(n)),
               "jpeggpu_b200_get_scan_stats")
        return n.value

    def scan_stats(self, scan=0):
        st = ScanStats()
        _check(self.lib.jpeggpu_b200_get_scan_stats(self.handle, scan, C.byref(st), None),
               "jpeggpu_b200_get_scan_stats")
        return st

    def transfer_size(self):
        n = C.c_size_t()
        _check(self.lib.jpeggpu_b200_get_transfer_size(self.handle, C.byref(n)),
               "jpeggpu_b200_get_transfer_size")
        return n.value

    def read_scan_info(self, d_tmp, scan=0, stream=0):
        """What the device found in the entropy-coded data of the last decode (synchronises)."""
        si = ScanInfo()
        _check(self.lib.jpeggpu_b200_read_scan_info(self.handle, scan, d_tmp, stream, C.byref(si)),
               "jpeggpu_b200_read_scan_info")
        return si

    def set_deferred_walk(self, on):
        """Who walks the entropy-coded bytes: 0 host at parse time, 1/2 host inside transfer,
        3 the device (default); include/jpeggpu_b200.h."""
        _check(self.lib.jpeggpu_b200_set_deferred_walk(self.handle, int(on)), "jpeggpu_b200_set_deferred_walk")

    def set_profiling(self, on):
        _check(self.lib.jpeggpu_b200_set_profiling(self.handle, int(on)), "jpeggpu_b200_set_profiling")

    def kernel_times(self):
        """{kernel name: (milliseconds, launches)} of the last decode (stream must be synced)."""
        ms, n = (C.c_float * MAX_KERNELS)(), (C.c_int * MAX_KERNELS)()
        _check(self.lib.jpeggpu_b200_get_kernel_times(self.handle, C.byref(ms), C.byref(n)),
               "jpeggpu_b200_get_kernel_times")
        out = {}
        for i in range(MAX_KERNELS):
            name = self.lib.jpeggpu_b200_kernel_name(i).decode()
            if name and n[i]:
                out[name] = (ms[i], n[i])
        return out

    def cleanup(self):
        if self.handle:
            _check(self.lib.jpeggpu_decoder_cleanup(self.handle), "jpeggpu_decoder_cleanup")
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.cleanup()
        except Exception:
            pass
