"""BASELINE config 3: non-interleaved and mixed scan structures (the files Pillow / stb cannot
write are produced by tests/jpeg_transcode.py from the coefficient planes of an interleaved
file, so the expected samples are known: those of the source file)."""
import numpy as np
import pytest

import emul_bindings as eb
import inputs
import jpeg_transcode as jt
import oracle_bindings as ob

CASES = {
    # name: (source kwargs, scans, restart interval, DHT before every SOS)
    "444_three_scans": (dict(w=203, h=131, seed=31, mode="YCbCr", subsampling=0), [[0], [1], [2]], 0, True),
    "420_three_scans": (dict(w=203, h=131, seed=32, mode="YCbCr", subsampling=2), [[0], [1], [2]], 0, True),
    "420_y_then_cbcr": (dict(w=203, h=131, seed=33, mode="YCbCr", subsampling=2), [[0], [1, 2]], 0, True),
    "422_three_scans_dri": (dict(w=200, h=120, seed=34, mode="YCbCr", subsampling=1), [[0], [1], [2]], 5, True),
    "cmyk_four_scans": (dict(w=160, h=96, seed=35, mode="CMYK"), [[0], [1], [2], [3]], 0, True),
    "444_tables_once": (dict(w=96, h=64, seed=36, mode="YCbCr", subsampling=0), [[0], [1], [2]], 0, False),
    "420_reordered": (dict(w=120, h=88, seed=37, mode="YCbCr", subsampling=2), [[2], [0, 1]], 3, True),
    "gray_single": (dict(w=77, h=50, seed=38, mode="L"), [[0]], 4, True),
    # the restart interval redefined between the scans (T.81 B.2.4.4; the reference answers
    # NOT_SUPPORTED, src/reader.cpp:563-568 - SURVEY 8f rank 4), switched off again for the last scan
    "420_dri_per_scan": (dict(w=203, h=131, seed=39, mode="YCbCr", subsampling=2), [[0], [1], [2]], [7, 3, 0], True),
    "444_dri_grows": (dict(w=160, h=120, seed=40, mode="YCbCr", subsampling=0), [[0], [1, 2]], [2, 50], True),
}


def build(name):
    kw, scans, ri, every = CASES[name]
    src = inputs.pillow_jpeg(**kw)
    ref = ob.decode(src)
    assert ref["status"] == 0
    data = jt.transcode(src, ref["planes"], scans, restart_interval=ri, dht_every_scan=every)
    return src, ref, data


@pytest.mark.parametrize("name", list(CASES))
def test_oracle_decodes_transcoded_file_to_the_source_samples(name):
    src, ref, data = build(name)
    out = ob.decode(data)
    assert out["status"] == 0
    assert out["info"].num_scans == len(CASES[name][1])
    for c in range(ref["info"].num_components):
        assert np.array_equal(out["pixels"][c], ref["pixels"][c]), f"component {c}"
        h, w = out["planes"][c].shape
        hh, ww = min(h, ref["planes"][c].shape[0]), min(w, ref["planes"][c].shape[1])
        assert np.array_equal(out["planes"][c][:hh, :ww], ref["planes"][c][:hh, :ww])


@pytest.mark.parametrize("name", list(CASES))
def test_emulated_pipeline_on_transcoded_file(name):
    _, _, data = build(name)
    out = ob.decode(data)
    for s in range(out["info"].num_scans):
        st, coef, stats, _ = eb.decode_scan(data, s, out["stream"][s].size)
        assert st == 0
        assert np.array_equal(coef, out["stream"][s]), f"scan {s}"


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(CASES))
def test_gpu_decodes_transcoded_file(name):
    from jpeggpu_b200.torch_io import decode_file
    src, ref, data = build(name)
    out = ob.decode(data)
    info, planes, coefs = decode_file(data, want_coefficients=True)
    assert info.num_components == ref["info"].num_components
    for c in range(info.num_components):
        assert np.array_equal(planes[c], ref["pixels"][c]), f"{name}: samples of component {c}"
        assert np.array_equal(coefs[c], out["planes"][c]), f"{name}: coefficients of component {c}"


@pytest.mark.gpu
def test_reference_agrees_on_6mp_four_component_four_scan_file():
    import ref_bindings as rb
    if not rb.available():
        pytest.skip("oracle/_ref/libjpeggpu_ref.so not built")
    from jpeggpu_b200.torch_io import decode_file
    data = inputs.four_scan_6mp()
    orc = ob.decode(data)
    assert orc["status"] == 0 and orc["info"].num_scans == 4
    oi = orc["info"]
    info, planes, coefs = decode_file(data, want_coefficients=True)
    assert info.num_components == 4 and (info.sizes_x[0], info.sizes_y[0]) == (3008, 2000)
    cw = [c.shape[1] for c in coefs]
    ch = [c.shape[0] for c in coefs]
    comps = [[oi.scans[s].comp_idx[a] for a in range(oi.scans[s].num_components)] for s in range(oi.num_scans)]
    st, rinfo, rplanes, rcoefs = rb.decode_with_reference(
        data, [oi.scans[s].num_segments for s in range(oi.num_scans)], cw, ch, comps)
    assert st == 0, st
    for c in range(4):
        assert np.array_equal(rplanes[c], orc["pixels"][c]), f"oracle samples differ from reference, comp {c}"
        assert np.array_equal(rplanes[c], planes[c]), f"product samples differ from reference, comp {c}"
        if rcoefs is not None:
            assert np.array_equal(rcoefs[c], coefs[c]), f"product coefficients differ from reference, comp {c}"


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["444_three_scans", "cmyk_four_scans"])
def test_reference_agrees_on_non_interleaved_1x1_scans(name):
    """The reference handles non-interleaved scans only for 1x1 sampling and only when every
    scan carries its own DHT; on those files all three decoders agree."""
    import ref_bindings as rb
    if not rb.available():
        pytest.skip("oracle/_ref/libjpeggpu_ref.so not built")
    from jpeggpu_b200.torch_io import decode_file
    _, _, data = build(name)
    orc = ob.decode(data)
    oi = orc["info"]
    info, planes, coefs = decode_file(data, want_coefficients=True)
    cw = [c.shape[1] for c in coefs] + [0] * (4 - len(coefs))
    ch = [c.shape[0] for c in coefs] + [0] * (4 - len(coefs))
    comps = [[oi.scans[s].comp_idx[a] for a in range(oi.scans[s].num_components)]
             for s in range(oi.num_scans)]
    st, rinfo, rplanes, rcoefs = rb.decode_with_reference(
        data, [oi.scans[s].num_segments for s in range(oi.num_scans)], cw, ch, comps)
    assert st == 0, st
    for c in range(info.num_components):
        assert np.array_equal(rplanes[c], planes[c]), f"{name}: samples differ from reference, comp {c}"
        assert np.array_equal(rplanes[c], orc["pixels"][c])
        if rcoefs is not None:
            assert np.array_equal(rcoefs[c], coefs[c]), f"{name}: coefficients differ from reference, comp {c}"
