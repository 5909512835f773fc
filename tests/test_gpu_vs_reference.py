"""Pins both the oracle and the product against the REFERENCE's own CUDA decoder
(oracle/_ref/libjpeggpu_ref.so, compiled from /root/reference/src by oracle/Makefile), run on
the same GPU on the same files: quantised coefficients bit-exact, samples bit-exact."""
import numpy as np
import pytest

import inputs
import oracle_bindings as ob
import ref_bindings as rb

pytestmark = pytest.mark.gpu


def _cases():
    out = []
    d = inputs.img6510()
    if d is not None:
        out.append(("IMG_6510", d))
    # (8064x6048: BASELINE config 4; amplitude 8: its > 16 MiB twin)
    for (w, h, seed, q, amp) in [(4032, 3024, 6510, 90, 1.0), (3008, 2000, 1, 90, 1.0), (3008, 2000, 2, 95, 1.0),
                                 (3001, 1999, 3, 90, 1.0), (8064, 6048, 48, 90, 1.0), (8064, 6048, 48, 90, 8.0)]:
        d = inputs.synth(w, h, seed, q, amp)
        if d is not None:
            out.append((f"stb{w}x{h}q{q}" + ("" if amp == 1.0 else f"a{amp:g}"), d))
    out.append(("pil422", inputs.pillow_jpeg(w=3008, h=2000, seed=12, mode="YCbCr", subsampling=1)))
    out.append(("pilgray", inputs.pillow_jpeg(w=3008, h=2000, seed=14, mode="L")))
    out.append(("pilrst", inputs.pillow_jpeg(w=1008, h=752, seed=17, mode="YCbCr", subsampling=2,
                                             restart_blocks=63)))
    return out


@pytest.mark.parametrize("label,data", _cases(), ids=lambda v: v if isinstance(v, str) else "")
def test_reference_cuda_decoder_agrees(label, data):
    if not rb.available():
        pytest.skip("oracle/_ref/libjpeggpu_ref.so not built")
    from jpeggpu_b200.torch_io import decode_file
    orc = ob.decode(data)
    assert orc["status"] == 0
    oi = orc["info"]
    info, planes, coefs = decode_file(data, want_coefficients=True)
    cw = [c.shape[1] for c in coefs] + [0] * (4 - len(coefs))
    ch = [c.shape[0] for c in coefs] + [0] * (4 - len(coefs))
    comps = [[oi.scans[s].comp_idx[a] for a in range(oi.scans[s].num_components)]
             for s in range(oi.num_scans)]
    st, rinfo, rplanes, rcoefs = rb.decode_with_reference(
        data, [oi.scans[s].num_segments for s in range(oi.num_scans)], cw, ch, comps)
    assert st == 0, st
    assert rinfo.num_components == info.num_components
    for c in range(info.num_components):
        assert (rinfo.sizes_x[c], rinfo.sizes_y[c]) == (info.sizes_x[c], info.sizes_y[c])
        assert (rinfo.subsampling.x[c], rinfo.subsampling.y[c]) == (info.subsampling.x[c], info.subsampling.y[c])
        # reference vs oracle (pins the oracle), reference vs product (the parity claim)
        assert np.array_equal(rplanes[c], orc["pixels"][c]), f"{label}: oracle samples differ from reference, comp {c}"
        assert np.array_equal(rplanes[c], planes[c]), f"{label}: product samples differ from reference, comp {c}"
        if rcoefs is not None:
            assert np.array_equal(rcoefs[c], orc["planes"][c]), f"{label}: oracle coefficients differ from reference, comp {c}"
            assert np.array_equal(rcoefs[c], coefs[c]), f"{label}: product coefficients differ from reference, comp {c}"
