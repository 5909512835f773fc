"""ctypes bindings of the REFERENCE decoder built from /root/reference sources into
oracle/_ref/libjpeggpu_ref.so (oracle/Makefile).  TEST INFRASTRUCTURE: used on the GPU box to pin
the oracle and the product against the reference's own CUDA implementation."""
import ctypes as C
import os

import numpy as np

from jpeggpu_b200 import Img, ImgInfo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libjpeggpu_ref.so")


def available():
    return os.path.exists(REF_LIB)


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(REF_LIB)
        dec = C.c_void_p
        L.jpeggpu_decoder_startup.argtypes = [C.POINTER(dec)]
        L.jpeggpu_decoder_parse_header.argtypes = [dec, C.POINTER(ImgInfo), C.c_void_p, C.c_size_t]
        L.jpeggpu_decoder_get_buffer_size.argtypes = [dec, C.POINTER(C.c_size_t)]
        L.jpeggpu_decoder_transfer.argtypes = [dec, C.c_void_p, C.c_size_t, C.c_void_p]
        L.jpeggpu_decoder_decode.argtypes = [dec, C.POINTER(Img), C.c_void_p, C.c_size_t, C.c_void_p]
        L.jpeggpu_decoder_cleanup.argtypes = [dec]
        _lib = L
    return _lib


def ceil256(v):
    return (v + 255) // 256 * 256


class RefDecoder:
    def __init__(self):
        self.L = lib()
        self.h = C.c_void_p()
        st = self.L.jpeggpu_decoder_startup(C.byref(self.h))
        assert st == 0, st

    def parse(self, ptr, size):
        info = ImgInfo()
        st = self.L.jpeggpu_decoder_parse_header(self.h, C.byref(info), ptr, size)
        return st, info

    def buffer_size(self):
        n = C.c_size_t()
        st = self.L.jpeggpu_decoder_get_buffer_size(self.h, C.byref(n))
        assert st == 0, st
        return n.value

    def transfer(self, d_tmp, size, stream=0):
        return self.L.jpeggpu_decoder_transfer(self.h, d_tmp, size, stream)

    def decode(self, planes, pitches, d_tmp, size, stream=0):
        img = Img()
        for c, (p, pitch) in enumerate(zip(planes, pitches)):
            img.image[c] = p
            img.pitch[c] = pitch
        return self.L.jpeggpu_decoder_decode(self.h, C.byref(img), d_tmp, size, stream)

    def close(self):
        if self.h:
            self.L.jpeggpu_decoder_cleanup(self.h)
            self.h = C.c_void_p()


def decode_with_reference(data: bytes, num_segments_per_scan, coef_w, coef_h, scan_components):
    """Run the reference decoder on the current CUDA device.  Returns (status, info, pixel planes,
    coefficient planes or None).  Coefficients are read back from the reference's scratch at the
    deterministic offsets of SURVEY.md §9.3 (single-scan files only; src/decoder.cpp:120-155,
    :240-246)."""
    import torch
    host = torch.empty(len(data), dtype=torch.uint8, pin_memory=True)
    host.numpy()[:] = np.frombuffer(data, dtype=np.uint8)
    rd = RefDecoder()
    st, info = rd.parse(host.data_ptr(), len(data))
    if st != 0:
        rd.close()
        return st, info, None, None
    need = rd.buffer_size()
    tmp = torch.zeros(need, dtype=torch.uint8, device="cuda")
    planes = [torch.zeros((info.sizes_y[c], info.sizes_x[c]), dtype=torch.uint8, device="cuda")
              for c in range(info.num_components)]
    st = rd.transfer(tmp.data_ptr(), need, 0)
    assert st == 0, st
    st = rd.decode([p.data_ptr() for p in planes], [p.stride(0) for p in planes], tmp.data_ptr(),
                   need, 0)
    torch.cuda.synchronize()
    rd.close()
    if st != 0:
        return st, info, None, None
    coefs = None
    if len(num_segments_per_scan) == 1:
        off = ceil256(len(data)) + 4 * 256 + 4 * ceil256(8 * 896)
        off += ceil256(8 * num_segments_per_scan[0])
        coefs = [None] * info.num_components
        for c in scan_components[0]:
            n = coef_w[c] * coef_h[c]
            raw = tmp[off:off + 2 * n].cpu().numpy().view(np.int16).reshape(coef_h[c], coef_w[c])
            coefs[c] = raw.copy()
            off += ceil256(2 * n)
    return 0, info, [p.cpu().numpy() for p in planes], coefs
