"""GPU parity suite of the plain `path` integrator (the metric's equal-time comparator): the CUDA
path through the C ABI against golden fixtures produced by the REAL reference (tests/golden/make_golden_path.py).
Tolerances: FP64 mode within 1e-9 relative of the reference's double arithmetic per sample (differences: libm last-ulp);
FP32 mode close on most samples; occlusion and light ids exact."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu

PATH_SCENES = ["cbox_gdpt_64", "cbox_gdpt_d2", "veach_mi_small", "sponza_small", "matpreview_small", "disney_small"]   # the last two: environment map


@pytest.fixture(scope="module")
def scenes(g):
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = g.Scene.named(name)
        return cache[name]
    yield get
    for s in cache.values():
        s.close()


@pytest.mark.parametrize("name", PATH_SCENES)
def test_light_sampling_matches_reference(g, scenes, name):
    z = golden("lights_%s.npz" % name)
    got = g.sample_light(scenes(name), z["ref_points"], z["rnd"], precision="f64")
    want = z["out"]
    assert np.array_equal(got[:, 0], want[:, 0])                            # sample_light: TableDist1D inversion
    assert np.allclose(got[:, 1:7], want[:, 1:7], rtol=1e-9, atol=1e-9)     # point + normal
    assert np.allclose(got[:, 7], want[:, 7], rtol=1e-8)                    # pdf_point_on_light
    assert np.allclose(got[:, 8], want[:, 8], rtol=1e-12)                   # light_pmf
    got32 = g.sample_light(scenes(name), z["ref_points"], z["rnd"], precision="f32")
    same = got32[:, 0] == want[:, 0]
    assert same.mean() > 0.995                                              # a float cdf may flip a boundary case
    assert np.allclose(got32[same, 1:7], want[same, 1:7], rtol=2e-3, atol=2e-3)


@pytest.mark.parametrize("name", PATH_SCENES)
def test_occluded_matches_reference(g, scenes, name):
    z = golden("lights_%s.npz" % name)
    got = g.occluded(scenes(name), z["occ_rays"], precision="f64")
    bad = int((got != z["occluded"]).sum())
    assert bad <= 2, "%d of %d occlusion queries differ" % (bad, got.size)   # grazing shadow rays: tfar rounding at the light


@pytest.mark.parametrize("name", PATH_SCENES)
def test_path_tracing_records_match_reference(g, scenes, name):
    z = golden("path_%s.npz" % name)
    got = g.path_samples(scenes(name), z["xs"], z["ys"], z["streams"], precision="f64")
    want = z["rad"]
    ok = np.isfinite(want).all(axis=1)
    bad = (np.abs(got[ok] - want[ok]) > 1e-9 * (1 + np.abs(want[ok]))).any(axis=1)
    assert bad.sum() <= max(2, 0.002 * ok.sum()), "%d of %d path_tracing samples differ" % (bad.sum(), ok.sum())


@pytest.mark.parametrize("name", ["cbox_gdpt_64", "veach_mi_small", "disney_small", "matpreview_small"])
def test_path_fp32_records_close(g, scenes, name):
    """matpreview_small is the hard one for float: rough glass at the roughness clamp (GGX a2 = 1e-8) lit by an
    environment map only."""
    z = golden("path_%s.npz" % name)
    got = g.path_samples(scenes(name), z["xs"], z["ys"], z["streams"], precision="f32")
    want = z["rad"]
    assert np.isfinite(got).all(), "%d float samples are not finite" % (~np.isfinite(got).all(axis=1)).sum()
    close = (np.abs(got - want) <= 2e-3 * (1 + np.abs(want))).all(axis=1)
    assert close.mean() > 0.85                       # float paths diverge at branch decisions (RR, occlusion)
    assert abs(got.mean() - want.mean()) < 0.05 * want.mean()


def test_path_pixel_seeded_image_matches_reference(g, scenes):
    want = golden("path_image_cbox64.npz")["img"]
    sc = scenes("cbox_gdpt_64")
    got = g.path_render(sc, spp=4, precision="f64")
    bad = (np.abs(got - want) > 1e-9 * (1 + np.abs(want))).any(axis=2)
    assert bad.mean() < 0.002, "%d pixels differ" % bad.sum()
    # render() dispatches on the integrator like render.cpp:373-395
    saved = sc.options.integrator
    sc.options.integrator = g.Integrator.Path
    try:
        assert np.allclose(g.render(sc, spp=4, precision="f64"), got, rtol=1e-12, atol=1e-15)   # atomic accumulation order
    finally:
        sc.options.integrator = saved


def test_path_image_mean_matches_the_reference_estimate(g, scenes):
    """64 fresh spp of the f32 bench mode against the mean of the reference's 3000 golden samples (statistical)."""
    sc = scenes("cbox_gdpt_64")
    p = g.path_render(sc, spp=64, precision="f32", sample0=1000)
    want = golden("path_cbox_gdpt_64.npz")["rad"].mean()
    assert abs(p.mean() - want) < 0.06 * want


def test_path_full_size_sponza(g, scenes):
    sc = scenes("sponza_1080p")
    img = g.path_render(sc, spp=1, precision="f32")
    assert np.isfinite(img).all() and img.mean() > 0.005
    st = g.stats(sc)
    assert st["rays_traced"] > 1920 * 1080 * 3


def test_path_refuses_an_environment_map_that_was_not_provided(g, scenes):
    """Containers written before the envmap export carry the light row but not its image/transforms: path refuses, gradpath
    (which ignores environment maps as shipped) still renders."""
    a = {k: v for k, v in scenes("disney_small").a.items() if not k.startswith("envmap.")}
    sc = g.Scene(a)
    try:
        with pytest.raises(g._capi.GdptError):
            g.path_render(sc, spp=1)
        assert np.isfinite(g.gradient_path_render(sc, spp=1, precision="f32").primal).all()
    finally:
        sc.close()


def test_path_environment_map_background(g, scenes):
    """Primary rays that leave the scene return the environment map (path_tracing.h:28-38): with max_depth = 1 (no loop
    iteration) the image is the background where the pixel sees no geometry and black elsewhere (no emitter in view)."""
    sc = scenes("matpreview_small")
    a = g.path_render(sc, spp=1, precision="f64", max_depth=1)
    b = g.path_render(sc, spp=1, precision="f64", max_depth=1)
    assert np.array_equal(a, b) and np.isfinite(a).all()
    ys, xs = np.mgrid[0:sc.height, 0:sc.width]
    sp = np.stack([(xs.ravel() + 0.5) / sc.width, (ys.ravel() + 0.5) / sc.height], axis=1)
    rays6 = g.sample_primary(sc, sp)
    rays = np.concatenate([rays6, np.zeros((sp.shape[0], 1)), np.full((sp.shape[0], 1), np.inf)], axis=1)
    ids, _ = g.intersect(sc, rays)
    miss = (ids[:, 0] < 0).reshape(sc.height, sc.width)
    lit = (a.sum(axis=2) > 0)
    if miss.any():
        assert lit[miss].mean() > 0.9          # the sample is jittered inside the pixel: silhouette pixels may differ
    assert lit[~miss].mean() < 0.1


@pytest.mark.parametrize("name", ["cbox_gdpt_64", "veach_mi_small"])
def test_path_queues_and_tail_kernel_agree(g, name, monkeypatch):
    """Per-bounce queue kernels (k_trace closest + any-hit, k_path_shade; with and without ray sorting) and the
    single-launch k_path_tail are the same arithmetic: FP64 records agree bit for bit with each other and the reference."""
    z = golden("path_%s.npz" % name)
    out = {}
    for tag, env in (("tail", {"GDPT_TAIL_PATHS": str(1 << 30)}), ("queues", {"GDPT_TAIL_PATHS": "0"})):
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        sc = g.Scene.named(name)
        out[tag] = g.path_samples(sc, z["xs"], z["ys"], z["streams"], precision="f64")
        sc.close()
    assert np.array_equal(out["tail"], out["queues"])
    bad = (np.abs(out["queues"] - z["rad"]) > 1e-9 * (1 + np.abs(z["rad"]))).any(axis=1)
    assert bad.sum() <= 6
    imgs = {}
    for tag, env in (("sorted", {"GDPT_TAIL_PATHS": "0", "GDPT_SORT_RAYS": "1"}), ("plain", {"GDPT_TAIL_PATHS": "0", "GDPT_SORT_RAYS": "0"}),
                     ("tail", {"GDPT_TAIL_PATHS": str(1 << 30), "GDPT_SORT_RAYS": "0"})):
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        sc = g.Scene.named("cbox_gdpt")
        imgs[tag] = g.path_render(sc, spp=1, precision="f64")
        sc.close()
    for tag in ("plain", "tail"):
        assert np.allclose(imgs["sorted"], imgs[tag], rtol=1e-9, atol=1e-12), tag


@pytest.mark.parametrize("name", ["cbox_gdpt_64", "veach_mi_small", "sponza_small"])
def test_aux_integrators_match_reference(g, scenes, name):
    """aux_render (render.cpp:15-72): depth, shading normal, mean curvature, ray differential and mip level of the primary
    vertex at every pixel centre, against the reference's own render() for Integrator values 0..4."""
    z = golden("aux_images.npz")
    sc = scenes(name)
    for kind in range(5):
        want = z["%s.%d" % (name, kind)]
        got = g.aux_render(sc, integrator=kind, precision="f64")
        ok = np.isclose(got, want, rtol=1e-9, atol=1e-12, equal_nan=True).all(axis=2)
        assert ok.mean() > 0.998, "integrator %d: %d pixels differ" % (kind, (~ok).sum())   # pixel-centre rays on a silhouette edge
    saved = sc.options.integrator
    sc.options.integrator = g.Integrator.Depth
    try:
        assert np.array_equal(g.render(sc), g.aux_render(sc, integrator=0))
    finally:
        sc.options.integrator = saved
    depth32 = g.aux_render(sc, integrator=0, precision="f32")
    assert np.allclose(depth32, z["%s.0" % name], rtol=1e-3, atol=1e-3) or (np.isclose(depth32, z["%s.0" % name], rtol=1e-3, atol=1e-3).mean() > 0.99)
