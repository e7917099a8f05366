"""TEST INFRASTRUCTURE — not part of the product.

L1 reconstruction of a gradient-domain render by iteratively reweighted least squares (Kettunen et al. 2015, "Gradient-Domain
Path Tracing", section 5.2 — the paper the reference's handout cites, handouts/final_project.tex:109-111).  The reference's
code has NO counterpart: both of its programs reconstruct with the L2 screened-Poisson solve only (fourierSolve,
src/render.cpp:172-254, small_gdpt.cpp:355-436), so parity for this function is UNPINNED; this file is the CPU statement the
GPU solver (csrc/gdpt_poisson_l1.cu) is tested against, written with exact sparse solves where the GPU iterates.

Energy, for an image I of [h][w][3] (gx(x, y) estimates I(x, y) - I(x-1, y) for x >= 1, gy(x, y) estimates I(x, y) - I(x, y-1) for
y >= 1 — the merged gradients of render.cpp:340-350; column 0 of gx and row 0 of gy are not used):

    E(I) = alpha * sum_p wU_p |I_p - u_p|^2 + sum_{x>=1} wX_p |I_p - I_(p-x) - gx_p|^2 + sum_{y>=1} wY_p |I_p - I_(p-y) - gy_p|^2

With all weights 1 this is the L2 problem.  IRLS: wX, wY = 1 / (eps + r), r = the L1 norm over the three channels of the term's
current residual (one weight per pixel and term, shared by the channels), eps = epsilon * mean |u|; wU = 1 by default (only the
gradient terms are made robust, so the image keeps the primal's mean exactly: sum_p of D^T W (D I - g) telescopes to zero), or
reweighted the same way with reweight_primal (the all-L1 energy of the paper); start from the L2 solution.
"""
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla


def _operators(w, h):
    n = w * h
    idx = np.arange(n).reshape(h, w)
    # Dx: rows for x >= 1 (row index = pixel index, rows of column 0 are empty), same for Dy
    px = idx[:, 1:].ravel(); qx = idx[:, :-1].ravel()
    Dx = sp.csr_matrix((np.r_[np.ones(px.size), -np.ones(px.size)], (np.r_[px, px], np.r_[px, qx])), shape=(n, n))
    py = idx[1:, :].ravel(); qy = idx[:-1, :].ravel()
    Dy = sp.csr_matrix((np.r_[np.ones(py.size), -np.ones(py.size)], (np.r_[py, py], np.r_[py, qy])), shape=(n, n))
    mx = np.zeros(n); mx[px] = 1.0
    my = np.zeros(n); my[py] = 1.0
    return Dx, Dy, mx, my


def weighted_solve(u, gx, gy, alpha, wU, wX, wY, ops=None):
    """argmin of the weighted energy (exact sparse solve per channel).  u, gx, gy: [h][w][3]; weights: [h][w]."""
    h, w, _ = u.shape
    Dx, Dy, mx, my = ops or _operators(w, h)
    WU = sp.diags(alpha * wU.ravel()); WX = sp.diags(wX.ravel() * mx); WY = sp.diags(wY.ravel() * my)
    A = (WU + Dx.T @ WX @ Dx + Dy.T @ WY @ Dy).tocsc()
    lu = spla.splu(A)
    out = np.empty_like(u)
    for c in range(3):
        b = WU @ u[..., c].ravel() + Dx.T @ (WX @ gx[..., c].ravel()) + Dy.T @ (WY @ gy[..., c].ravel())
        out[..., c] = lu.solve(b).reshape(h, w)
    return out


def weights(I, u, gx, gy, eps, reweight_primal=False):
    h, w, _ = u.shape
    rU = np.abs(I - u).sum(-1) if reweight_primal else np.full((h, w), 1.0 - eps)      # 1 / (eps + 1 - eps) = 1
    rX = np.zeros((h, w)); rX[:, 1:] = np.abs(I[:, 1:] - I[:, :-1] - gx[:, 1:]).sum(-1)
    rY = np.zeros((h, w)); rY[1:] = np.abs(I[1:] - I[:-1] - gy[1:]).sum(-1)
    return 1.0 / (eps + rU), 1.0 / (eps + rX), 1.0 / (eps + rY)


def reconstruct_l1(u, gx, gy, alpha=0.04, irls_iterations=10, epsilon=0.05, start=None, reweight_primal=False):
    """IRLS from `start` (default: the all-ones-weights solution).  Returns the image after `irls_iterations` reweightings.
    reweight_primal=False keeps wU = 1 (the primal term stays quadratic: the image keeps the primal's mean exactly);
    True reweights it like the gradient terms (the all-L1 energy of the paper, which pulls towards per-pixel medians)."""
    u = np.asarray(u, np.float64); gx = np.asarray(gx, np.float64); gy = np.asarray(gy, np.float64)
    h, w, _ = u.shape
    ops = _operators(w, h)
    one = np.ones((h, w))
    I = weighted_solve(u, gx, gy, alpha, one, one, one, ops) if start is None else np.array(start, np.float64)
    eps = epsilon * max(float(np.abs(u).mean()), 1e-300)
    for _ in range(irls_iterations):
        wU, wX, wY = weights(I, u, gx, gy, eps, reweight_primal)
        I = weighted_solve(u, gx, gy, alpha, wU, wX, wY, ops)
    return I


def energy_l1(I, u, gx, gy, alpha, reweight_primal=False):
    """The objective the IRLS iterations decrease (per-pixel L1 norms over the channels; the primal term quadratic unless
    reweight_primal)."""
    e = alpha * (np.abs(I - u).sum() if reweight_primal else ((I - u) ** 2).sum())
    e += np.abs(I[:, 1:] - I[:, :-1] - gx[:, 1:]).sum()
    e += np.abs(I[1:] - I[:-1] - gy[1:]).sum()
    return float(e)
