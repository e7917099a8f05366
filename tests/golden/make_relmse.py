"""Golden data for the convergence gate (north star: "primal/gradient images must converge to the
reference's 16k-spp render with relMSE matching the CPU path at equal spp within 5 %").

Run in the build container (needs oracle/_ref):  python tests/golden/make_relmse.py
Writes tests/golden/relmse_cbox64.npz:
  R_primal, R_dx, R_dy, R_poisson   the compiled reference at 16384 spp on cbox_gdpt_64 (pixel-seeded streams)
  seeds                             K RNG seeds
  cpu_relmse[K][4]                  relMSE of the reference at SPP spp (one row per seed) against R
  spp
relMSE = mean((I - R)^2 / (R^2 + 1e-3))  (SURVEY.md 8d)."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import cport, ref  # noqa: E402

SPP_REF, SPP, K = 16384, 16, 12


def relmse(i, r):
    return float(np.mean((i - r) ** 2 / (r ** 2 + 1e-3)))


def images(s, spp, seed):
    c, cx0, cy0, cx1, cy1 = s.render_pixelseed(spp, seed=seed)
    cx, cy = cport.merge(cx0, cx1, cy0, cy1)
    return c, cx, cy, cport.poisson(c, cx, cy)


def main():
    s = ref.RefScene(ref.scene_path("cbox_gdpt_64"))
    R = images(s, SPP_REF, 0x9E3779B97F4A7C15)
    seeds = np.array([1000003 * (k + 1) + 17 for k in range(K)], dtype=np.uint64)
    cpu = np.array([[relmse(i, r) for i, r in zip(images(s, SPP, int(sd)), R)] for sd in seeds])
    s.close()
    np.savez_compressed(os.path.join(HERE, "relmse_cbox64.npz"), R_primal=R[0], R_dx=R[1], R_dy=R[2], R_poisson=R[3],
                        seeds=seeds, cpu_relmse=cpu, spp=SPP)
    print("cpu relMSE mean over seeds (primal, dx, dy, poisson):", cpu.mean(axis=0))


if __name__ == "__main__":
    main()
