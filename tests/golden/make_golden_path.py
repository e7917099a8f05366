#!/usr/bin/env python
"""Golden fixtures of the plain `path` integrator, produced by the REAL reference
(oracle/_ref/liblajolla_ref.so, built by oracle/build_ref.sh from /root/reference/src).  Run in the build
container; the .npz files are committed so the GPU box (no /root/reference) can check parity.

  path_<scene>.npz    path_tracing (src/path_tracing.h:13-348) radiance for seeded (x, y, stream) triples,
                      rng = init_pcg32(stream)
  lights_<scene>.npz  sample_light + sample_point_on_light + pdf_point_on_light + light_pmf
                      (src/scene.cpp:61-67, src/light.cpp:57-71) and occluded (src/intersection.cpp:66-85)
  path_image_cbox64.npz  pixel-seeded 4-spp image (stream = sample*W*H + y*W + x)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402


def main():
    ref.stage_scenes()
    rng = np.random.default_rng(20261105)
    only = sys.argv[1:]
    for name, n in [("cbox_gdpt_64", 3000), ("cbox_gdpt_d2", 1500), ("veach_mi_small", 3000), ("sponza_small", 3000),
                    ("matpreview_small", 3000), ("disney_small", 3000)]:      # the last two: environment map (+ a sphere light)
        if only and name not in only:
            continue
        s = ref.RefScene(ref.scene_path(name))
        xs = rng.integers(0, s.width, n).astype(np.int32); ys = rng.integers(0, s.height, n).astype(np.int32)
        st = rng.integers(0, 2 ** 40, n).astype(np.uint64)
        rad = np.array([s.path_sample(int(x), int(y), int(t)) for x, y, t in zip(xs, ys, st)])
        np.savez_compressed(os.path.join(HERE, "path_%s.npz" % name), xs=xs, ys=ys, streams=st, rad=rad,
                            max_depth=np.array(s.max_depth), rr_depth=np.array(s.rr_depth))
        # light sampling seen from surface points (primary hits) and from random points in space
        m = 2000
        sp = rng.random((m, 2))
        rays6 = s.sample_primary(sp)
        rays = np.concatenate([rays6, np.zeros((m, 1)), np.full((m, 1), np.inf)], axis=1)
        ids, vtx = s.intersect(rays, None)
        pts = vtx[ids[:, 0] >= 0, 0:3]
        rnd = rng.random((pts.shape[0], 4))
        out = s.sample_light(pts, rnd)
        # shadow rays towards the sampled points, as path_tracing builds them (:131-134) ...
        d = out[:, 1:4] - pts
        dist = np.linalg.norm(d, axis=1, keepdims=True)
        eps = 1e-4
        env = (out[:, 1:4] == 0).all(axis=1)          # environment-map samples: direction = -normal, shadow ray to infinity
        d[env] = -out[env, 4:7]; dist[env] = 1.0
        tfar = (1 - eps) * dist
        tfar[env] = np.inf
        shadow = np.concatenate([pts, d / dist, np.full_like(dist, eps), tfar], axis=1)
        # ... and segments between random pairs of surface points
        perm = rng.permutation(pts.shape[0])
        d2 = pts[perm] - pts
        dist2 = np.linalg.norm(d2, axis=1, keepdims=True)
        ok = dist2[:, 0] > 1e-6
        seg = np.concatenate([pts[ok], d2[ok] / dist2[ok], np.full_like(dist2[ok], eps), (1 - eps) * dist2[ok]], axis=1)
        occ_rays = np.concatenate([shadow, seg])
        np.savez_compressed(os.path.join(HERE, "lights_%s.npz" % name), ref_points=pts, rnd=rnd, out=out,
                            occ_rays=occ_rays, occluded=s.occluded(occ_rays))
        s.close()
    # auxiliary integrators (aux_render, src/render.cpp:15-72): depth, shading normal, mean curvature, ray differential, mip level
    if not only or "aux" in only:
        out = {}
        for name in ("cbox_gdpt_64", "veach_mi_small", "sponza_small"):
            s = ref.RefScene(ref.scene_path(name))
            for kind in range(5):
                out["%s.%d" % (name, kind)] = s.render_aux(kind)
            s.close()
        np.savez_compressed(os.path.join(HERE, "aux_images.npz"), **out)
    if not only:
        s = ref.RefScene(ref.scene_path("cbox_gdpt_64"))
        np.savez_compressed(os.path.join(HERE, "path_image_cbox64.npz"), img=s.render_path_pixelseed(4))
        s.close()
    print("path golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
