"""oracle/poisson_l1.py — the CPU statement of the L1 (IRLS) reconstruction the GPU solver is compared with.  The reference's
code has no L1 reconstruction (parity unpinned); what CAN be pinned here is that the statement's L2 limit is the reference's
fourierSolve up to its border convention, and that the reweighting does what it is for."""
import numpy as np


def synthetic(h=48, w=64, seed=1, outliers=0.02, primal_noise=0.2):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    I0 = np.stack([0.5 + 0.4 * np.sin(x / 9.0) * np.cos(y / 7.0), (x > w // 2) * 0.6 + 0.2, 0.3 + 0.01 * y], -1)
    gx = np.zeros_like(I0); gx[:, 1:] = I0[:, 1:] - I0[:, :-1]
    gy = np.zeros_like(I0); gy[1:] = I0[1:] - I0[:-1]
    u = I0 + rng.normal(0, primal_noise, I0.shape)
    gxn = gx + rng.normal(0, 0.01, gx.shape); gyn = gy + rng.normal(0, 0.01, gy.shape)
    if outliers > 0:
        m = rng.random((h, w)) < outliers; gxn[m] += rng.normal(0, 3, (int(m.sum()), 3))
        m = rng.random((h, w)) < outliers; gyn[m] += rng.normal(0, 3, (int(m.sum()), 3))
    return I0, u, gx, gy, gxn, gyn


def test_l2_limit_is_fouriersolve_up_to_the_border_convention():
    """All weights 1, consistent gradients: the statement's minimiser against the oracle's fourierSolve fed with the border
    entries that function expects (gx(0, y) = -gx(1, y): its whole-sample symmetric extension, render.cpp:216-224)."""
    from oracle import cport, poisson_l1 as L
    I0, u, gx, gy, _, _ = synthetic(primal_noise=0.05, outliers=0)
    mine = L.reconstruct_l1(u, gx, gy, 0.04, irls_iterations=0)
    gxf = gx.copy(); gxf[:, 0] = -gx[:, 1]
    gyf = gy.copy(); gyf[0] = -gy[1]
    theirs = cport.poisson(u, gxf, gyf, 0.04)
    assert np.linalg.norm(mine - theirs) / np.linalg.norm(theirs) < 2e-3
    inner = (slice(6, -6), slice(6, -6))
    assert np.linalg.norm((mine - theirs)[inner]) / np.linalg.norm(theirs[inner]) < 5e-4


def test_reweighting_rejects_gradient_outliers_and_lowers_the_l1_energy():
    from oracle import poisson_l1 as L
    I0, u, _, _, gxn, gyn = synthetic()
    l2 = L.reconstruct_l1(u, gxn, gyn, 0.04, 0)
    l1 = L.reconstruct_l1(u, gxn, gyn, 0.04, 10)
    rmse = lambda a: float(np.sqrt(((a - I0) ** 2).mean()))
    assert rmse(l1) < 0.25 * rmse(l2), (rmse(l1), rmse(l2))
    e = [L.energy_l1(L.reconstruct_l1(u, gxn, gyn, 0.04, k), u, gxn, gyn, 0.04) for k in (0, 1, 3, 10)]
    assert e[0] > e[1] > e[2] > e[3], e
    # quadratic primal term: the mean of the primal is kept exactly; the all-L1 energy does not keep it
    assert abs(l1.mean() - u.mean()) < 1e-10
    full = L.reconstruct_l1(u, gxn, gyn, 0.04, 10, reweight_primal=True)
    assert rmse(full) < 0.25 * rmse(l2) and abs(full.mean() - u.mean()) > 1e-6
