"""B200 timing of gdpt_poisson_l1 at film sizes (device milliseconds from the entry point's own events)."""
import sys
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
import gradient_based_path_tracing_b200 as g
from test_poisson_l1_cpu import synthetic
for (h, w) in ((512, 512), (1080, 1920)):
    I0, u, _, _, gxn, gyn = synthetic(h=h, w=w, seed=2)
    rmse = lambda a: float(np.sqrt(((a - I0) ** 2).mean()))
    l2 = g.fourier_solve(u, gxn, gyn, 0.04)
    for irls, cg in ((10, 50), (10, 200), (20, 100)):
        out, info = g.reconstruct_l1(u, gxn, gyn, 0.04, irls_iterations=irls, cg_iterations=cg, return_info=True)
        print("%4dx%-4d  irls %2d x cg <= %3d: %8.1f ms device, %5d cg steps taken, last relative residual %.1e;  rmse vs truth: L1 %.4f, fourierSolve %.4f, primal %.4f" % (
            w, h, irls, cg, info["ms"], info["cg_steps"], info["relative_residual"], rmse(out), rmse(l2), rmse(u)), flush=True)
