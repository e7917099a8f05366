#!/usr/bin/env python
"""Round-2 golden fixtures, produced by the REAL reference (oracle/_ref) in the build container:

  filter_<scene>.npz   Box(2) / Tent(2) / Gaussian(2) pixel filters (src/filters/*.inl; the three filters of
                       src/tests/filter.cpp:15-66) on variants of the 64x64 Cornell box: sample_primary rays for
                       seeded screen positions + grad_path_tracing records (5 filtered primaries per sample)
  config_<scene>.npz   BASELINE configs C1-C3 at FULL film size (cbox_gdpt 512x512, veach_mi_gradpath 768x512,
                       disney_gradpath 683x512): the reference's pixel-seeded accumulation at SPP_CFG spp, stored as
                       8x8-block sums of the five images (a checksum of checksums: every pixel enters one sum) plus
                       the image means
  relmse_<scene>.npz   the convergence gate on the small films of Sponza / veach_mi / Disney (tests/golden/make_relmse.py
                       does the Cornell box): 16384-spp reference images + relMSE of the CPU path at 16 spp, 12 seeds

usage: make_golden_r2.py [filters] [configs] [relmse[:scene]]"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from oracle import cport, ref  # noqa: E402

FILTER_SCENES = ["cbox_gdpt_64_box", "cbox_gdpt_64_tent", "cbox_gdpt_64_gauss2"]
CONFIG_SCENES = {"cbox_gdpt": 2, "veach_mi_gradpath": 1, "disney_gradpath": 1}      # name -> spp
RELMSE_SCENES = ["veach_mi_small", "sponza_small", "disney_small"]
SPP_REF, SPP, K = 16384, 16, 12
BLOCK = 8


def block_sums(img):
    h, w, c = img.shape
    hp, wp = -(-h // BLOCK) * BLOCK, -(-w // BLOCK) * BLOCK
    pad = np.zeros((hp, wp, c)); pad[:h, :w] = img
    return pad.reshape(hp // BLOCK, BLOCK, wp // BLOCK, BLOCK, c).sum(axis=(1, 3))


def filters():
    rng = np.random.default_rng(20261105)
    for name in FILTER_SCENES:
        s = ref.RefScene(ref.scene_path(name))
        sp = rng.random((4000, 2)) * 1.04 - 0.02
        # the inputs of src/tests/filter.cpp: rnd_param = (0.3, 0.4) and its two finite-difference neighbours, as the
        # fractional part of a pixel position (camera.cpp:23-47 feeds the fraction to sample(filter, .))
        eps = 1e-6
        px, py = 20, 30
        kat = np.array([[(px + 0.3) / s.width, (py + 0.4) / s.height], [(px + 0.3 + eps) / s.width, (py + 0.4) / s.height],
                        [(px + 0.3) / s.width, (py + 0.4 + eps) / s.height]])
        sp = np.concatenate([kat, sp])
        rays = s.sample_primary(sp)
        n = 1500
        xs = rng.integers(0, s.width, n).astype(np.int32); ys = rng.integers(0, s.height, n).astype(np.int32)
        st = rng.integers(0, 2 ** 40, n).astype(np.uint64)
        rec = np.array([s.grad_sample(int(x), int(y), int(t)) for x, y, t in zip(xs, ys, st)])
        np.savez_compressed(os.path.join(HERE, "filter_%s.npz" % name), screen_pos=sp, rays=rays, xs=xs, ys=ys, streams=st, rec=rec)
        s.close()
        print(name, "done")


def configs():
    for name, spp in CONFIG_SCENES.items():
        t0 = time.time()
        s = ref.RefScene(ref.scene_path(name))
        imgs = s.render_pixelseed(spp)
        cx, cy = cport.merge(imgs[1], imgs[3], imgs[2], imgs[4])
        out = {"spp": np.array(spp), "size": np.array([s.width, s.height])}
        for key, im in zip(("c", "cx", "cy"), [imgs[0], cx, cy]):      # = primal, dx, dy of gdpt_render (render.cpp:340-350)
            im = np.nan_to_num(im, nan=0.0, posinf=0.0, neginf=0.0)
            out["bs_" + key] = block_sums(im); out["abs_" + key] = block_sums(np.abs(im)); out["mean_" + key] = im.mean(axis=(0, 1))
        out["nonfinite"] = np.array(sum(int((~np.isfinite(im)).sum()) for im in imgs))
        np.savez_compressed(os.path.join(HERE, "config_%s.npz" % name), **out)
        s.close()
        print(name, "%.1f s" % (time.time() - t0), "primal mean", out["mean_c"], "non-finite", out["nonfinite"])


def relmse(i, r):
    """mean((I - R)^2 / (R^2 + 1e-3)) over the values finite in both images (the reference's Disney lobes produce NaN
    records — pow of negative bases — which poison single pixels of its own images)."""
    ok = np.isfinite(i) & np.isfinite(r)
    return float(np.mean((i[ok] - r[ok]) ** 2 / (r[ok] ** 2 + 1e-3)))


def images(s, spp, seed):
    c, cx0, cy0, cx1, cy1 = s.render_pixelseed(spp, seed=seed)
    cx, cy = cport.merge(cx0, cx1, cy0, cy1)
    return c, cx, cy, cport.poisson(c, cx, cy)


def relmse_scene(name):
    t0 = time.time()
    s = ref.RefScene(ref.scene_path(name))
    R = images(s, SPP_REF, 0x9E3779B97F4A7C15)
    seeds = np.array([1000003 * (k + 1) + 17 for k in range(K)], dtype=np.uint64)
    cpu = np.array([[relmse(i, r) for i, r in zip(images(s, SPP, int(sd)), R)] for sd in seeds])
    s.close()
    np.savez_compressed(os.path.join(HERE, "relmse_%s.npz" % name), R_primal=R[0], R_dx=R[1], R_dy=R[2], R_poisson=R[3],
                        seeds=seeds, cpu_relmse=cpu, spp=SPP)
    print(name, "%.0f s" % (time.time() - t0), "cpu relMSE mean over seeds (primal, dx, dy, poisson):", cpu.mean(axis=0))


def main():
    what = sys.argv[1:] or ["filters", "configs", "relmse"]
    ref.stage_scenes()
    for w in what:
        if w == "filters":
            filters()
        elif w == "configs":
            configs()
        elif w.startswith("relmse"):
            for name in ([w.split(":")[1]] if ":" in w else RELMSE_SCENES):
                relmse_scene(name)


if __name__ == "__main__":
    main()
