"""Shift-mapped gradpath (gdpt_render_params.offset_mode = GDPT_OFFSETS_SHIFT; csrc/gdpt_shift.cuh): the algorithm of
small_gdpt.cpp:153-352 on lajolla scenes.  No reference implementation runs it on these scenes (small_gdpt.cpp has its own
sphere scene, camera and RNG; lajolla's gradpath is the as-shipped variant), so there are no reference outputs to pin it
to.  The records are compared with a CPU statement of the algorithm on the reference's own primitives (last test), and
the rest are the properties a gradient-domain estimator must have: gradients that converge to the finite differences of the
converged primal, a Poisson image that converges to the primal and beats it at equal sample count, the border
convention of the screened-Poisson solve, determinism, band composition."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def g():
    import gradient_based_path_tracing_b200 as g
    return g


@pytest.fixture(scope="module")
def converged(g):
    sc = g.Scene.named("cbox_gdpt_64")
    return sc, g.gradient_path_render(sc, spp=32768, precision="f32", offsets="shift", sample0=100000)


def rel_l2(a, b):
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))


def relmse(a, ref):
    return float(np.mean((a - ref) ** 2 / (ref ** 2 + 1e-3)))


def test_gradients_converge_to_the_finite_differences_of_the_primal(g, converged):
    sc, T = converged
    Dx = T.primal[:, 1:] - T.primal[:, :-1]
    Dy = T.primal[1:] - T.primal[:-1]
    # measured on B200: 0.019 / 0.0075 at 65536 spp, 0.0096 / 0.0038 at 262144 spp (noise of the primal's differences)
    assert rel_l2(T.dx[:, 1:], Dx) < 0.06
    assert rel_l2(T.dy[1:], Dy) < 0.03
    # the as-shipped estimator does not have this property (offsets are dropped after the first bounce): 0.6 / 0.87
    R = g.gradient_path_render(sc, spp=4096, precision="f32")
    assert rel_l2(R.dx[:, 1:], Dx) > 0.3 and rel_l2(R.dy[1:], Dy) > 0.3


def test_border_columns_follow_the_solver_convention(converged):
    """fourierSolve reads gx(0, y) as u(0, y) - u(1, y) (render.cpp:216-224): column 0 / row 0 hold minus column 1 / row 1."""
    _, T = converged
    assert np.allclose(T.dx[:, 0], -T.dx[:, 1], rtol=1e-9, atol=1e-14)
    assert np.allclose(T.dy[0], -T.dy[1], rtol=1e-9, atol=1e-14)


def test_poisson_image_converges_to_the_primal_and_beats_it(g, converged):
    sc, T = converged
    assert rel_l2(T.poisson, T.primal) < 0.01                       # measured 0.0039 at 65536 spp
    for spp in (16, 64):
        a = g.gradient_path_render(sc, spp=spp, precision="f32", offsets="shift")
        # measured: 0.77 vs 2.71 at 16 spp, 0.20 vs 0.74 at 64 spp
        assert relmse(a.poisson, T.primal) < 0.5 * relmse(a.primal, T.primal)


def test_primal_agrees_with_the_reference_mode(g, converged):
    """Both modes estimate the same image with BSDF sampling only (different roulette, 24-vertex cap here)."""
    sc, T = converged
    R = g.gradient_path_render(sc, spp=8192, precision="f32", sample0=200000, outputs=("primal",))
    assert abs(T.primal.mean() - R.primal.mean()) < 0.03 * R.primal.mean()   # measured 1.0 %


def test_records_are_consistent_and_deterministic(g):
    sc = g.Scene.named("veach_mi_small")
    rng = np.random.default_rng(3)
    n = 4000
    xs, ys = rng.integers(0, sc.width, n), rng.integers(0, sc.height, n)
    st = np.arange(n, dtype=np.uint64) + 77
    a = g.grad_samples(sc, xs, ys, st, precision="f64", offsets="shift")
    b = g.grad_samples(sc, xs, ys, st, precision="f64", offsets="shift")
    assert np.array_equal(a, b)
    rad, contrib, prob, w = a[:, 0:3], a[:, 3:6], a[:, 18], a[:, 19:23]
    assert np.isfinite(a).all() and (prob > 0).all()
    assert ((w > 0) & (w <= 1)).all()
    assert np.allclose(rad, contrib / prob[:, None], rtol=1e-12, atol=0)
    hit = (contrib > 0).any(axis=1)
    assert 0.02 < hit.mean() < 0.9                                   # base paths that ended on a light
    cs = a[:, 6:18].reshape(n, 4, 3)
    assert (cs >= 0).all() and (cs[~hit] == 0).all()                 # no offsets without a base path
    shifted = (cs[hit] > 0).any(axis=2)
    assert shifted.mean() > 0.3                                      # most shifts of a successful base succeed
    f32 = g.grad_samples(sc, xs, ys, st, precision="f32", offsets="shift")
    close = np.abs(f32[:, 0:3] - rad) <= 2e-3 * (1e-3 + np.abs(rad))
    assert close.all(axis=1).mean() > 0.9


def test_row_bands_compose_to_the_full_image(g):
    sc = g.Scene.named("cbox_gdpt_64")
    full = g.gradient_path_render(sc, spp=3, precision="f64", offsets="shift", outputs=("primal", "dx", "dy"))
    # bands as the multi-GPU split makes them (multiples of 4 rows): the five accumulation images add up; dx is band-local
    acc_p, acc_dx = np.zeros_like(full.primal), np.zeros_like(full.dx)
    for y0, y1 in ((0, 20), (20, 44), (44, 64)):
        part = g.gradient_path_render(sc, spp=3, precision="f64", offsets="shift", rows=(y0, y1), outputs=("primal", "dx"))
        acc_p += part.primal
        acc_dx[y0:y1] += part.dx[y0:y1]
    assert np.allclose(acc_p, full.primal, rtol=1e-12, atol=1e-15)
    assert np.allclose(acc_dx, full.dx, rtol=1e-9, atol=1e-13)
    # the merge is linear, so whole merged images add up too — also for a band edge between rows 0 and 1, where the border rule
    # of row 0 (minus the row-1 estimate) has to be carried by the band that owns row 1
    acc = {n: np.zeros_like(full.primal) for n in ("primal", "dx", "dy")}
    for y0, y1 in ((0, 1), (1, 20), (20, 44), (44, 64)):
        part = g.gradient_path_render(sc, spp=3, precision="f64", offsets="shift", rows=(y0, y1), outputs=("primal", "dx", "dy"))
        for n in acc:
            acc[n] += part[n]
    for n in acc:
        assert np.allclose(acc[n], full[n], rtol=1e-9, atol=1e-13), n


def test_glossy_and_textured_scenes_render(g):
    for name, lo in (("veach_mi_small", 0.0), ("sponza_small", 0.0)):
        sc = g.Scene.named(name)
        r = g.gradient_path_render(sc, spp=16, precision="f32", offsets="shift")
        for im in (r.primal, r.dx, r.dy, r.poisson):
            assert np.isfinite(im).all()
        assert r.primal.mean() > lo
        ref = g.gradient_path_render(sc, spp=16, precision="f32", outputs=("primal",))
        assert r.stats["rays_traced"] > 0 and ref.primal.mean() >= 0


def test_unknown_offset_mode_is_rejected(g):
    import ctypes as C
    from gradient_based_path_tracing_b200 import _capi
    from gradient_based_path_tracing_b200.render import _params
    sc = g.Scene.named("cbox_gdpt_64")
    p = _params(sc, spp=1)
    p.offset_mode = 7
    imgs = _capi.GdptImages()
    assert _capi.lib().gdpt_render(sc.context(0), C.byref(p), C.byref(imgs)) != 0
    assert b"offset mode" in _capi.lib().gdpt_last_error()


def test_band_renderer_carries_the_offset_mode(g):
    """multi.DistributedRenderer (the bench's and the multi-GPU path's front end) with one rank equals gdpt_render."""
    import torch
    from gradient_based_path_tracing_b200 import multi

    class Solo:
        ReduceOp = type("R", (), {"SUM": 0})
        def get_rank(self): return 0
        def get_world_size(self): return 1
        def reduce(self, *a, **k): pass
        def barrier(self): pass

    sc = g.Scene.named("cbox_gdpt_64")
    want = g.gradient_path_render(sc, spp=4, precision="f64", offsets="shift")
    R = multi.DistributedRenderer(sc, 0, Solo(), torch)
    got = R.render(4, precision="f64", to_host=True, offsets="shift")
    for k in ("primal", "dx", "dy", "poisson"):
        assert np.allclose(got[k], getattr(want, k), rtol=1e-9, atol=1e-13), k
    ref = R.render(4, precision="f64", to_host=True)
    assert not np.allclose(ref["dx"], got["dx"])


def test_state_machine_kernel_equals_the_straight_one(g, monkeypatch):
    """k_shift_sm (per-lane phases, one ray-cast site, lanes fetch their next sample) against k_shift (program order):
    same arithmetic in the same order, so FP64 records are bit-identical; images agree up to the order of the atomics."""
    out, img = {}, {}
    rng = np.random.default_rng(5)
    for name in ("veach_mi_small", "sponza_small"):
        sc0 = g.Scene.named(name)
        n = 3000
        xs, ys = rng.integers(0, sc0.width, n), rng.integers(0, sc0.height, n)
        st = np.arange(n, dtype=np.uint64) + 11
        for tag in ("0", "1"):
            monkeypatch.setenv("GDPT_SHIFT_KERNEL", tag)
            out[tag] = g.grad_samples(sc0, xs, ys, st, precision="f64", offsets="shift")
        assert np.array_equal(out["0"], out["1"]), name
        assert name != "veach_mi_small" or (out["1"][:, 3:6] > 0).any()
    sc = g.Scene.named("cbox_gdpt_64")
    for tag in ("0", "1"):
        monkeypatch.setenv("GDPT_SHIFT_KERNEL", tag)
        img[tag] = g.gradient_path_render(sc, spp=8, precision="f64", offsets="shift")
    for k in ("primal", "dx", "dy", "poisson"):
        assert np.allclose(getattr(img["0"], k), getattr(img["1"], k), rtol=1e-9, atol=1e-13), k


@pytest.mark.parametrize("name", ["cbox_gdpt_64", "cbox_gdpt_d2", "veach_mi_small", "disney_small"])
def test_records_match_the_cpu_statement_on_the_reference_primitives(g, name):
    """FP64 records against tests/golden/shift_<scene>.npz: the same algorithm written on the reference's own
    sample_primary / intersect / occluded / sample_bsdf / eval / pdf / emission (oracle/ref_shims/ref_hooks.cpp,
    gdpt_ref_shift_sample; fixtures by tests/golden/make_golden_shift.py).  Lambertian, rough plastic, Disney BSDFs,
    sphere and mesh lights, a depth-2 scene; both kernels."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "shift_%s.npz" % name))
    sc = g.Scene.named(name)
    want = z["rec"]
    hit = (want[:, 3:6] > 0).any(axis=1)
    assert hit.sum() >= 50
    for kernel in ("1", "0"):
        os.environ["GDPT_SHIFT_KERNEL"] = kernel
        try:
            got = g.grad_samples(sc, z["xs"], z["ys"], z["streams"], precision="f64", offsets="shift")
        finally:
            del os.environ["GDPT_SHIFT_KERNEL"]
        bad = (np.abs(got - want) > 1e-9 * (1e-300 + np.abs(want))).any(axis=1)
        # a last-ulp difference may flip a branch (roulette, lobe choice, a grazing shadow ray): a handful of samples at most
        assert bad.sum() <= max(2, 0.002 * len(want)), "%s kernel %s: %d of %d records differ (%d of them among the %d base paths that reached a light)" % (
            name, kernel, bad.sum(), len(want), (bad & hit).sum(), hit.sum())


def test_image_pipeline_equals_the_oracle_records_accumulated_in_numpy(g):
    """The whole image path of the mode — accumulation signs, the y pairing, the border rule, the merge, the screened
    Poisson solve — against the CPU statement: every (pixel, sample) record from the oracle, accumulated here in numpy
    the way finish_shift / k_merge do it, solved with the oracle's fourierSolve restatement."""
    from oracle import ref, cport
    if not (ref.available() and ref.scenes_available()):
        pytest.skip("oracle/_ref not built")
    sc = g.Scene.named("cbox_gdpt_64")
    W, H, spp = sc.width, sc.height, 2
    rs = ref.RefScene(ref.scene_path("cbox_gdpt_64"), threads=1)
    img = np.zeros((H, W, 3)); cx0 = np.zeros((H, W, 3)); cy0 = np.zeros((H, W, 3)); cx1 = np.zeros((H, W, 3)); cy1 = np.zeros((H, W, 3))
    try:
        for s in range(spp):
            for y in range(H):
                for x in range(W):
                    r = rs.shift_sample(x, y, s * W * H + y * W + x)
                    c, p = r[3:6], r[18]
                    img[y, x] += r[0:3] / spp
                    vx0 = (c - r[6:9]) * (r[19] / (p * spp)); vx1 = (r[9:12] - c) * (r[21] / (p * spp))
                    vy0 = (c - r[12:15]) * (r[20] / (p * spp)); vy1 = (r[15:18] - c) * (r[22] / (p * spp))
                    if x > 0:
                        cx0[y, x] += vx0
                        if x == 1:
                            cx0[y, 0] -= vx0
                    if y > 0:
                        cy0[y, x] += vy0
                        if y == 1:
                            cy0[0, x] -= vy0
                    if x < W - 1:
                        cx1[y, x] += vx1
                        if x == 0:
                            cx0[y, 0] -= vx1
                    if y < H - 1:
                        cy1[y, x] += vy1
                        if y == 0:
                            cy0[0, x] -= vy1
    finally:
        rs.close()
    cx = cx0.copy(); cx[:, 1:] += cx1[:, :-1]
    cy = cy0.copy(); cy[1:] += cy1[:-1]
    want_poisson = cport.poisson(img, cx, cy, 0.04)
    got = g.gradient_path_render(sc, spp=spp, precision="f64", offsets="shift")
    assert img.any() and cx.any() and cy.any()
    for name, a, b in (("primal", got.primal, img), ("dx", got.dx, cx), ("dy", got.dy, cy)):
        close = np.abs(a - b) <= 1e-9 * (1e-12 + np.abs(b))
        assert close.all(axis=2).mean() > 0.998, "%s: %d pixels differ" % (name, (~close.all(axis=2)).sum())
    rel = np.linalg.norm(got.poisson - want_poisson) / np.linalg.norm(want_poisson)
    assert rel < 1e-5, rel
