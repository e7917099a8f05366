"""`./lajolla scene.xml` through the REAL reference parser and CLI with the GPU binding in the gradpath arm:
oracle/_ref/lajolla_gdpt = the reference's main.cpp + parse_scene + loaders + integration/render_b200.cpp + libgdpt.so
(integration/build_lajolla_gdpt.sh; built where /root/reference exists, travels with oracle/_ref/).  The XML, OBJ,
.serialized and JPG files are the staged copies of the reference's scene assets under oracle/_ref/scenes/."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

EXE = os.path.join(ROOT, "oracle", "_ref", "lajolla_gdpt")
SCENES = os.path.join(ROOT, "oracle", "_ref", "scenes")


def _read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = (int(v) for v in f.readline().split())
        f.readline()
        return np.frombuffer(f.read(), dtype="<f4").reshape(h, w, 3)


needs_exe = pytest.mark.skipif(not os.path.exists(EXE), reason="oracle/_ref/lajolla_gdpt not built (needs /root/reference at build time)")


@pytest.mark.gpu
@needs_exe
@pytest.mark.parametrize("xml,name,prec", [("cbox/gdpt_cbox_gdpt_64.xml", "cbox_gdpt_64", "f64"), ("cbox/gdpt_cbox_gdpt_64_tent.xml", "cbox_gdpt_64_tent", "f64"),
                                           ("veach_mi/gdpt_veach_mi_small.xml", "veach_mi_small", "f64"), ("sponza/gdpt_sponza_small.xml", "sponza_small", "f64"),
                                           ("sponza/gdpt_sponza_small.xml", "sponza_small", "f32")])
def test_reference_cli_with_gpu_gradpath_arm(g, tmp_path, xml, name, prec):
    """The reference's CLI parses the XML (its own parser, OBJ / .serialized / JPG loaders), render() dispatches
    Integrator::GradPath to gradient_path_render_b200, main() writes the Poisson image with the reference's imwrite.
    Result == the product's Python host on the pre-flattened container of the same scene (same C ABI underneath), and the
    primal / dx / dy buffers the binding can write beside it."""
    out = tmp_path / "img.pfm"
    env = dict(os.environ, GDPT_PRECISION=prec, GDPT_WRITE_BUFFERS=str(tmp_path / "buf"))
    r = subprocess.run([EXE, "-o", str(out), os.path.join(SCENES, xml)], capture_output=True, timeout=900, env=env, cwd=str(tmp_path))
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    text = r.stdout.decode()
    assert "Rendering..." in text and "Image written to" in text          # src/main.cpp:36-46
    sc = g.Scene.named(name)
    want = g.gradient_path_render(sc, precision=prec)                     # spp, depths: the scene's own (the XML's)
    sc.close()
    tol = dict(rtol=1e-6, atol=1e-9) if prec == "f64" else dict(rtol=1e-4, atol=1e-7)
    for path, img in ((out, want.poisson), (tmp_path / "buf_primal.pfm", want.primal), (tmp_path / "buf_dx.pfm", want.dx), (tmp_path / "buf_dy.pfm", want.dy)):
        got = _read_pfm(str(path))
        ok = np.isclose(got, img.astype(np.float32), **tol)
        assert ok.mean() > 0.999, (str(path), ok.mean())


@needs_exe
def test_reference_cli_binding_fails_loudly_without_a_gpu(g, tmp_path):
    """No CPU fallback behind the binding: without a device the gradpath arm throws (the reference's own error
    convention, src/flexception.h) instead of rendering on the host."""
    if g._capi.lib().gdpt_device_count() > 0:
        pytest.skip("a CUDA device is present")
    r = subprocess.run([EXE, "-o", str(tmp_path / "x.pfm"), os.path.join(SCENES, "cbox/gdpt_cbox_gdpt_64.xml")], capture_output=True, timeout=300)
    assert r.returncode != 0 and b"no CUDA device" in r.stderr
