"""BASELINE config 0: the reference's unmodified example program (example/example_tool.c, built
by oracle/Makefile once against the reference library and once against libjpeggpu_b200.so)
decodes images/IMG_6510.JPG; the PNGs it writes must be byte-identical."""
import hashlib
import os
import subprocess

import pytest

import gen_inputs

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_EXE = os.path.join(ROOT, "oracle", "_ref", "jpeggpu_example_ref")
OUR_EXE = os.path.join(ROOT, "oracle", "_ref", "jpeggpu_example_b200")


def _run(exe, src, dst):
    r = subprocess.run([exe, src, dst], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "gpu decode done" in r.stdout
    with open(dst, "rb") as f:
        return hashlib.sha256(f.read()).hexdigest()


def test_reference_example_tool_runs_unchanged_on_our_library(tmp_path):
    if not (os.path.exists(REF_EXE) and os.path.exists(OUR_EXE)):
        pytest.skip("example binaries not built (oracle/Makefile needs /root/reference)")
    src = gen_inputs.ensure_img6510()
    if src is None:
        pytest.skip("IMG_6510.JPG not staged")
    ours = _run(OUR_EXE, src, str(tmp_path / "ours.png"))
    ref = _run(REF_EXE, src, str(tmp_path / "ref.png"))
    assert ours == ref
