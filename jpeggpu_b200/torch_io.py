"""Device-memory plumbing for tests and bench.py: torch owns the pinned host buffer, the scratch
buffer, the output planes and the stream; every byte of decode work happens inside
libjpeggpu_b200.so.  Nothing here computes on the CPU."""
import numpy as np
import torch

from . import Decoder


def pinned_bytes(data: bytes) -> torch.Tensor:
    t = torch.empty(len(data), dtype=torch.uint8, pin_memory=True)
    t.numpy()[:] = np.frombuffer(data, dtype=np.uint8)
    return t


class DecodeSession:
    """Reusable decoder + scratch for one image at a time on the current CUDA device."""

    def __init__(self, device="cuda:0"):
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        self.dec = Decoder()
        self.scratch = None
        self.host = None
        self.info = None

    def parse(self, host: torch.Tensor):
        """host: pinned uint8 tensor holding the file."""
        self.host = host
        self.info = self.dec.parse_header(host.data_ptr(), host.numel(), keepalive=host)
        need = self.dec.get_buffer_size()
        if self.scratch is None or self.scratch.numel() < need:
            self.scratch = torch.empty(need + need // 4, dtype=torch.uint8, device=self.device)
        return self.info

    def alloc_planes(self, pitch_align=1):
        planes = []
        for c in range(self.info.num_components):
            w, h = self.info.sizes_x[c], self.info.sizes_y[c]
            pitch = (w + pitch_align - 1) // pitch_align * pitch_align
            planes.append(torch.zeros((h, pitch), dtype=torch.uint8, device=self.device))
        return planes

    def transfer(self, stream=None):
        s = stream.cuda_stream if stream is not None else torch.cuda.current_stream().cuda_stream
        self.dec.transfer(self.scratch.data_ptr(), self.scratch.numel(), s)

    def decode(self, planes, stream=None):
        s = stream.cuda_stream if stream is not None else torch.cuda.current_stream().cuda_stream
        self.dec.decode([p.data_ptr() for p in planes], [p.stride(0) for p in planes],
                        self.scratch.data_ptr(), self.scratch.numel(), s)

    def decode_rgbi(self, rgb=None, stream=None):
        """Fused decode to an (H, W, 3) uint8 device tensor (allocated when not given; any row
        stride >= 3*W is accepted)."""
        s = stream.cuda_stream if stream is not None else torch.cuda.current_stream().cuda_stream
        if rgb is None:
            rgb = torch.empty((self.info.sizes_y[0], self.info.sizes_x[0], 3), dtype=torch.uint8,
                              device=self.device)
        self.dec.decode_rgbi(rgb.data_ptr(), rgb.stride(0), self.scratch.data_ptr(),
                             self.scratch.numel(), s)
        return rgb

    def decode_coefficients(self, stream=None):
        s = stream.cuda_stream if stream is not None else torch.cuda.current_stream().cuda_stream
        w, h = self.dec.coefficient_size()
        planes = [torch.zeros((h[c], w[c]), dtype=torch.int16, device=self.device)
                  for c in range(self.info.num_components)]
        self.dec.decode_coefficients([p.data_ptr() for p in planes], self.scratch.data_ptr(),
                                     self.scratch.numel(), s)
        return planes

    def scan_info(self, scan=0):
        """jpeggpu_b200_read_scan_info for the last decode on the current stream (synchronises)."""
        return self.dec.read_scan_info(self.scratch.data_ptr(), scan,
                                       torch.cuda.current_stream().cuda_stream)

    def read_scratch(self, offset, nbytes):
        return self.scratch[offset:offset + nbytes].cpu().numpy()

    def close(self):
        self.dec.cleanup()


def decode_file(data: bytes, device="cuda:0", want_coefficients=False):
    """Convenience: full protocol for one image; returns (info, [uint8 planes], [int16 planes])."""
    sess = DecodeSession(device)
    host = pinned_bytes(data)
    info = sess.parse(host)
    sess.transfer()
    planes = sess.alloc_planes()
    sess.decode(planes)
    coefs = None
    if want_coefficients:
        coefs = sess.decode_coefficients()
    torch.cuda.synchronize()
    out = [p[:, : info.sizes_x[c]].cpu().numpy() for c, p in enumerate(planes)]
    cf = [c.cpu().numpy() for c in coefs] if coefs is not None else None
    sess.close()
    return info, out, cf
