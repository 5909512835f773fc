"""Cross-check of the CUDA path against nvJPEG from the local toolkit (NVJPEG_OUTPUT_UNCHANGED:
native subsampled planes, no colour conversion — the same artefact our decoder produces).  nvJPEG
uses a different IDCT, so this is a tolerance check ("max-abs <= 2 pixel
cross-checks, never bit-exact"); the bit-exact gates are test_gpu_parity.py (oracle) and
test_gpu_vs_reference.py (the reference's own CUDA decoder)."""
import numpy as np
import pytest

import inputs
import nvjpeg_bindings as nb

pytestmark = pytest.mark.gpu

# the reference IDCT truncates to int16 between its two passes (src/idct.cu:44-47, :98-144), so a
# few samples land 2 LSB away from a float IDCT; the bound below was never exceeded on B200
MAX_ABS = 2
MAX_MEAN_ABS = 0.25


def nvjpeg_planes(data):
    import torch
    host = torch.empty(len(data), dtype=torch.uint8, pin_memory=True)
    host.numpy()[:] = np.frombuffer(data, dtype=np.uint8)
    nj = nb.NvJpeg("gpu_hybrid")
    try:
        ch, ws, hs = nj.info(host.data_ptr(), len(data))
        planes = [torch.zeros((hs[c], ws[c]), dtype=torch.uint8, device="cuda") for c in range(ch)]
        st = nj.decode(host.data_ptr(), len(data), [p.data_ptr() for p in planes],
                       [p.stride(0) for p in planes], torch.cuda.current_stream().cuda_stream, True)
        assert st == 0, f"nvjpeg status {st}"
        return [p.cpu().numpy() for p in planes]
    finally:
        nj.close()


CASES = [("stb 12MP 4:2:0", lambda: inputs.synth(4032, 3024, 6510, 90)),
         ("stb 6MP 4:2:0", lambda: inputs.synth(3008, 2000, 1, 90)),
         ("stb 2MP 4:2:0 (cropped rows)", lambda: inputs.synth(1920, 1080, 1, 90)),
         ("stb 6MP odd size", lambda: inputs.synth(3001, 1999, 3, 90)),
         ("IMG_6510 (DRI 252)", inputs.img6510)]


@pytest.mark.parametrize("label,loader", CASES, ids=[c[0] for c in CASES])
def test_samples_close_to_nvjpeg(label, loader):
    if not nb.available():
        pytest.skip("oracle/_ref/libnvjpeg_harness.so not built")
    data = loader()
    if data is None:
        pytest.skip("input not staged")
    from jpeggpu_b200.torch_io import decode_file
    info, planes, _ = decode_file(data, want_coefficients=False)
    theirs = nvjpeg_planes(data)
    assert len(theirs) == info.num_components
    for c in range(info.num_components):
        a = np.asarray(planes[c]).astype(np.int16)
        b = theirs[c].astype(np.int16)
        assert a.shape == b.shape, (label, c, a.shape, b.shape)
        d = np.abs(a - b)
        assert int(d.max()) <= MAX_ABS, f"{label} comp {c}: max abs {int(d.max())}"
        assert float(d.mean()) <= MAX_MEAN_ABS, f"{label} comp {c}: mean abs {float(d.mean()):.4f}"
