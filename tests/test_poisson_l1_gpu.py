"""gdpt_poisson_l1 (include/gdpt.h): L1 reconstruction by iteratively reweighted least squares on the GPU, against its CPU
statement oracle/poisson_l1.py (exact sparse solves).  The reference's code has no L1 reconstruction — parity is unpinned; the
tolerances are those of the conjugate-gradient solves (written in each test)."""
import numpy as np
import pytest

from test_poisson_l1_cpu import synthetic

pytestmark = pytest.mark.gpu


def rel(a, b):
    return float(np.linalg.norm(a - b) / np.linalg.norm(b))


def test_l2_problem_by_conjugate_gradients_equals_the_exact_sparse_solve(g):
    from oracle import poisson_l1 as L
    _, u, _, _, gxn, gyn = synthetic()
    want = L.reconstruct_l1(u, gxn, gyn, 0.04, irls_iterations=0)
    got, info = g.reconstruct_l1(u, gxn, gyn, 0.04, irls_iterations=0, cg_iterations=3000, cg_tolerance=1e-13, start="primal", return_info=True)
    assert info["relative_residual"] <= 1e-12 and 10 < info["cg_steps"] < 3000, info
    assert rel(got, want) < 1e-9, rel(got, want)


def test_irls_iterations_follow_the_cpu_statement(g):
    from oracle import poisson_l1 as L
    _, u, _, _, gxn, gyn = synthetic()
    start = L.reconstruct_l1(u, gxn, gyn, 0.04, irls_iterations=0)
    for k, full in ((1, False), (4, False), (3, True)):
        want = L.reconstruct_l1(u, gxn, gyn, 0.04, irls_iterations=k, epsilon=0.05, reweight_primal=full)
        got, info = g.reconstruct_l1(u, gxn, gyn, 0.04, irls_iterations=k, cg_iterations=20000, cg_tolerance=1e-13, epsilon=0.05, start=start,
                                     reweight_primal=full, return_info=True)
        assert info["relative_residual"] <= 1e-12, info
        # k exact solves against k solves to 1e-13 relative residual of systems whose weights span ~3 orders of magnitude
        assert rel(got, want) < 1e-7, (k, full, rel(got, want))


def test_l1_rejects_outliers_where_fouriersolve_spreads_them(g):
    from oracle import poisson_l1 as L
    I0, u, _, _, gxn, gyn = synthetic(h=96, w=128, seed=5)
    l2 = g.fourier_solve(u, gxn, gyn, 0.04)
    l1, info = g.reconstruct_l1(u, gxn, gyn, 0.04, irls_iterations=10, cg_iterations=200, return_info=True)
    rmse = lambda a: float(np.sqrt(((a - I0) ** 2).mean()))
    assert rmse(l1) < 0.25 * rmse(l2), (rmse(l1), rmse(l2))
    assert L.energy_l1(l1, u, gxn, gyn, 0.04) < 0.6 * L.energy_l1(l2, u, gxn, gyn, 0.04)
    full = g.reconstruct_l1(u, gxn, gyn, 0.04, irls_iterations=10, cg_iterations=200, reweight_primal=True)
    assert rmse(full) < 0.25 * rmse(l2)
    # bit-reproducible: fixed-order reductions, no atomics
    again = g.reconstruct_l1(u, gxn, gyn, 0.04, irls_iterations=10, cg_iterations=200)
    assert np.array_equal(l1, again)


def test_on_a_rendered_frame_l1_is_a_valid_reconstruction(g):
    """Shift-mapped gradients of the Cornell box at 16 spp: the L1 image keeps the primal's mean and is closer to the
    high-sample primal than the primal itself (not necessarily closer than L2: this scene has no outliers)."""
    sc = g.Scene.named("cbox_gdpt_64")
    ref = g.gradient_path_render(sc, spp=2048, precision="f32", seed=11, offsets="shift").primal
    r = g.gradient_path_render(sc, spp=16, precision="f32", seed=3, offsets="shift")
    l1 = g.reconstruct_l1(r.primal, r.dx, r.dy, 0.04, irls_iterations=8, cg_iterations=200)
    sc.close()
    relmse = lambda a: float((((a - ref) ** 2) / (ref ** 2 + 1e-2)).mean())
    assert abs(l1.mean() / r.primal.mean() - 1) < 1e-3          # quadratic primal term: the mean is kept (to the CG tolerance)
    assert relmse(l1) < 0.7 * relmse(r.primal), (relmse(l1), relmse(r.primal), relmse(r.poisson))
    # the all-L1 energy pulls towards per-pixel medians: on this estimator (most samples are zero) it darkens the image
    full = g.reconstruct_l1(r.primal, r.dx, r.dy, 0.04, irls_iterations=8, cg_iterations=200, reweight_primal=True)
    assert full.mean() < 0.9 * r.primal.mean()


def test_argument_errors(g):
    _, u, _, _, gxn, gyn = synthetic(h=8, w=8)
    with pytest.raises(g._capi.GdptError):
        g.reconstruct_l1(u, gxn, gyn, irls_iterations=-1)
    with pytest.raises(g._capi.GdptError):
        g.reconstruct_l1(u, gxn, gyn, cg_iterations=0)
    with pytest.raises(ValueError):
        g.reconstruct_l1(u, gxn, gyn[:4])
    with pytest.raises(ValueError):
        g.reconstruct_l1(u, gxn, gyn, start="zero")
