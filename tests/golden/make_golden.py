#!/usr/bin/env python
"""Generates the golden fixtures in this directory by running the REAL reference
(oracle/_ref/liblajolla_ref.so = /root/reference/src compiled by oracle/build_ref.sh).
Run it in the build container (needs /root/reference); the .npz files it writes are
committed so the GPU box, which has no /root/reference, can check parity against them.

  poisson.npz        fourierSolve (src/render.cpp:172-254) on seeded random c/cx/cy, 3 sizes
  records_<scene>.npz  grad_path_tracing (src/path_tracing.h:354-1050) records for seeded
                     (x, y, stream) triples: rng = init_pcg32(stream)
  intersect_<scene>.npz  sample_primary + intersect (camera.cpp:23-47, intersection.cpp:7-65)
  bsdf_zoo.npz       eval / pdf_sample_bsdf / sample_bsdf (material.cpp:90-119) for every material of
                     the material-zoo container (scenes/materials_zoo.gdpt.z, also written here)
  render_cbox64.npz  pixel-seeded accumulation of 2 spp (c, cx0, cy0, cx1, cy1) + the reference's
                     own tile-seeded render() statistics at 64 spp
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402
from gradient_based_path_tracing_b200.scene import read_container, write_container  # noqa: E402

ZOO = [  # (reference scene xml relative to scenes/, label)
    ("veach_mi/mi.xml", "veach"), ("matpreview/matpreview.xml", "matpreview"),
    ("disney_bsdf_test/disney_bsdf.xml", "disney_bsdf"), ("disney_bsdf_test/disney_diffuse.xml", "disney_diffuse"),
    ("disney_bsdf_test/disney_metal.xml", "disney_metal"), ("disney_bsdf_test/disney_glass.xml", "disney_glass"),
    ("disney_bsdf_test/disney_clearcoat.xml", "disney_clearcoat"), ("disney_bsdf_test/disney_sheen.xml", "disney_sheen"),
    ("cbox/cbox_gdpt.xml", "cbox"),
]


def unit(v):
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


def bsdf_inputs(rng, n):
    """random shading configurations: frame, geometric normal near the shading normal (either side), dirs on both hemispheres"""
    nrm = unit(rng.standard_normal((n, 3)))
    t = unit(np.cross(nrm, unit(rng.standard_normal((n, 3)))))
    b = np.cross(nrm, t)
    gn = unit(nrm + 0.2 * rng.standard_normal((n, 3)))
    flip = rng.random(n) < 0.15
    gn[flip] *= -1
    uv = rng.random((n, 2)) * 3 - 1
    fp = rng.random((n, 1)) * 0.01
    vtx = np.concatenate([gn, t, b, nrm, uv, fp], axis=1)
    din = unit(nrm * rng.random((n, 1)) * 1.5 + rng.standard_normal((n, 3)))
    dout = unit(rng.standard_normal((n, 3)))
    same = rng.random(n) < 0.6     # bias towards the reflection hemisphere so lobes are exercised
    sgn = np.sign(np.sum(din * nrm, axis=1))
    dout[same] = unit(dout[same] + 1.2 * (nrm[same] * sgn[same, None]))
    return vtx, np.concatenate([din, dout], axis=1), rng.random((n, 3))


def main():
    ref.stage_scenes()
    rng = np.random.default_rng(20261104)

    # ---- Poisson
    out = {}
    for i, (h, w) in enumerate([(17, 23), (64, 48), (96, 127)]):
        u = rng.random((h, w, 3)); gx = rng.normal(0, 0.1, (h, w, 3)); gy = rng.normal(0, 0.1, (h, w, 3))
        out.update({"u%d" % i: u, "gx%d" % i: gx, "gy%d" % i: gy, "out%d" % i: ref.fourier_solve(u, gx, gy, 0.04)})
    u = rng.random((31, 45, 3)); gx = rng.normal(0, 0.1, u.shape); gy = rng.normal(0, 0.1, u.shape)
    out.update({"u3": u, "gx3": gx, "gy3": gy, "out3": ref.fourier_solve(u, gx, gy, 0.4), "alpha3": np.array(0.4)})
    np.savez_compressed(os.path.join(HERE, "poisson.npz"), **out)

    # ---- per-sample records + intersections
    for name, n in [("cbox_gdpt_64", 3000), ("cbox_gdpt_d2", 1500), ("veach_mi_small", 3000), ("disney_small", 3000), ("sponza_small", 3000)]:
        s = ref.RefScene(ref.scene_path(name))
        xs = rng.integers(0, s.width, n).astype(np.int32); ys = rng.integers(0, s.height, n).astype(np.int32)
        st = rng.integers(0, 2 ** 40, n).astype(np.uint64)
        rec = np.array([s.grad_sample(int(x), int(y), int(t)) for x, y, t in zip(xs, ys, st)])
        np.savez_compressed(os.path.join(HERE, "records_%s.npz" % name), xs=xs, ys=ys, streams=st, rec=rec,
                            max_depth=np.array(s.max_depth), rr_depth=np.array(s.rr_depth))
        # primary rays through random screen positions (incl. slightly outside the film, as the offsets do)
        m = 4000
        sp = rng.random((m, 2)) * 1.04 - 0.02
        rays6 = s.sample_primary(sp)
        rays = np.concatenate([rays6, np.zeros((m, 1)), np.full((m, 1), np.inf)], axis=1)
        rd = np.tile(np.array([[0.0, 0.25 / max(s.width, s.height)]]), (m, 1))
        ids, vtx = s.intersect(rays, rd)
        # secondary rays: from hit points into random directions with the scene epsilon
        hit = ids[:, 0] >= 0
        org = vtx[hit, 0:3]
        k = org.shape[0]
        d2 = unit(rng.standard_normal((k, 3)))
        eps = 1e-3
        rays2 = np.concatenate([org, d2, np.full((k, 1), eps), np.full((k, 1), np.inf)], axis=1)
        ids2, vtx2 = s.intersect(rays2, None)
        np.savez_compressed(os.path.join(HERE, "intersect_%s.npz" % name), screen_pos=sp, rays=rays, rdiff=rd, ids=ids, vtx=vtx,
                            rays2=rays2, ids2=ids2, vtx2=vtx2)
        s.close()

    # ---- material zoo: materials of several reference scenes on a one-triangle scene
    mats, texs, labels = [], [], []
    zoo_scenes = []
    scenes_root = os.path.join(ref.STAGE_DIR)
    for rel, label in ZOO:
        s = ref.RefScene(os.path.join(scenes_root, rel))
        tmp = os.path.join(HERE, "_tmp.gdpt")
        s.export(tmp)
        a = read_container(tmp); os.remove(tmp)
        if any(k.startswith("img") and k.endswith(".size") for k in a) and label != "cbox":
            pass
        base = len(texs)
        for m in a["materials"]:
            m = m.copy()
            ids = m[2:]; ids[ids >= 0] += base
            mats.append(m); labels.append(label)
        for t in a["textures"]:
            if t[0] == 1:
                raise RuntimeError("zoo scenes must use constant/checker textures only")
            texs.append(t)
        zoo_scenes.append((s, len(mats) - len(a["materials"]), len(a["materials"])))
    base = read_container(os.path.join(ROOT, "scenes", "cbox_gdpt_64.gdpt.z"))
    zoo = {k: v for k, v in base.items() if not k.startswith("img")}
    zoo["materials"] = np.array(mats); zoo["textures"] = np.array(texs)
    write_container(os.path.join(ROOT, "scenes", "materials_zoo.gdpt.z"), zoo)
    gold = {"labels": np.array(labels)}
    for s, first, count in zoo_scenes:
        for j in range(count):
            vtx, dirs, rnd = bsdf_inputs(rng, 400)
            gold["vtx%d" % (first + j)] = vtx; gold["dirs%d" % (first + j)] = dirs; gold["rnd%d" % (first + j)] = rnd
            gold["out%d" % (first + j)] = s.bsdf(j, vtx, dirs, rnd)
        s.close()
    np.savez_compressed(os.path.join(HERE, "bsdf_zoo.npz"), **gold)

    # ---- image-level: pixel-seeded accumulation + tile-seeded render statistics
    s = ref.RefScene(ref.scene_path("cbox_gdpt_64"))
    c, cx0, cy0, cx1, cy1 = s.render_pixelseed(2)
    s.set_options(spp=64)
    r = s.render()
    np.savez_compressed(os.path.join(HERE, "render_cbox64.npz"), c=c, cx0=cx0, cy0=cy0, cx1=cx1, cy1=cy1,
                        ref64_primal=r["primal"], ref64_dx=r["dx"], ref64_dy=r["dy"], ref64_poisson=r["poisson"])
    s.close()
    print("golden fixtures written to", HERE)


if __name__ == "__main__":
    main()
