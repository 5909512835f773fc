"""Every device path once on small inputs - the command compute-sanitizer wraps
(tools/sanitize_gpu.sh): planes, coefficient planes, fused RGB, stand-alone RGB conversion, for
4:2:0 with restart markers, 4:4:4 with optimised tables, grey, a three-scan file and two damaged
files.  Results are checked against the oracle so that a sanitizer-clean run is also a correct one."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

import inputs  # noqa: E402
import jpeg_transcode as jt  # noqa: E402
import jpeggpu_b200 as jb  # noqa: E402
import oracle_bindings as ob  # noqa: E402
import test_gpu_robustness as rb  # noqa: E402
from jpeggpu_b200.torch_io import DecodeSession, pinned_bytes  # noqa: E402


def run(data, check=True):
    sess = DecodeSession("cuda:0")
    try:
        info = sess.parse(pinned_bytes(data))
        sess.transfer()
    except jb.JpegGpuError:
        sess.close()
        return "refused"
    planes = sess.alloc_planes()
    sess.decode(planes)
    coefs = sess.decode_coefficients()
    rgb = None
    if info.num_components in (1, 3):
        try:
            rgb = sess.decode_rgbi()
        except jb.JpegGpuError:
            rgb = torch.zeros((info.sizes_y[0], info.sizes_x[0], 3), dtype=torch.uint8, device="cuda:0")
            jb.convert_rgbi([p.data_ptr() for p in planes], [p.stride(0) for p in planes], info,
                            rgb.data_ptr(), rgb.stride(0), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    if check:
        ref = ob.decode(data)
        n = info.num_components
        for c in range(n):
            assert np.array_equal(planes[c][:, : info.sizes_x[c]].cpu().numpy(), ref["pixels"][c])
            assert np.array_equal(coefs[c].cpu().numpy(), ref["planes"][c])
        if rgb is not None:
            rc, want = ob.conv_to_rgbi(ref["pixels"], list(ref["info"].ssx)[:n], list(ref["info"].ssy)[:n])
            assert rc == 0 and np.array_equal(rgb.cpu().numpy(), want)
    sess.close()
    return "ok"


def main():
    a = inputs.pillow_jpeg(w=320, h=200, seed=71, mode="YCbCr", subsampling=2, restart_blocks=7)
    b = inputs.pillow_jpeg(w=203, h=117, seed=72, mode="YCbCr", subsampling=0, optimize=True, quality=95)
    c = inputs.pillow_jpeg(w=250, h=130, seed=73, mode="L")
    d = inputs.pillow_jpeg(w=320, h=200, seed=74, mode="YCbCr", subsampling=1)
    multi = jt.transcode(d, ob.decode(d)["planes"], scans=[[0], [1], [2]])
    for name, data in (("420_dri", a), ("444_opt", b), ("grey", c), ("422", d), ("three_scans", multi)):
        print(name, run(data))
    for kind in ("random_bytes", "markers", "shuffled", "truncated_eoi"):
        print("damaged", kind, run(rb.mutate(a, kind, 1), check=False))


if __name__ == "__main__":
    main()
