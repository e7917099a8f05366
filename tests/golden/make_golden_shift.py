#!/usr/bin/env python
"""Golden records of the shift-mapped gradpath mode (gdpt_render_params.offset_mode = GDPT_OFFSETS_SHIFT).

The reference has no such mode on lajolla scenes, so these are NOT outputs of reference code paths: they come from
oracle/ref_shims/ref_hooks.cpp::gdpt_ref_shift_sample, a CPU statement of the algorithm (small_gdpt.cpp:153-352) written
on the reference's own primitives (sample_primary, intersect, occluded, sample_bsdf / eval / pdf_sample_bsdf, emission)
and compiled into oracle/_ref/liblajolla_ref.so.  They pin the GPU kernels to an independent implementation that shares
every geometric and material routine with the reference.  Run in the build container; the .npz files are committed.

  shift_<scene>.npz   xs, ys, streams, rec[n][23]
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
    rng = np.random.default_rng(20261106)
    only = sys.argv[1:]
    for name, n in [("cbox_gdpt_64", 20000), ("cbox_gdpt_d2", 8000), ("veach_mi_small", 8000), ("disney_small", 10000)]:
        if only and name not in only:
            continue
        s = ref.RefScene(ref.scene_path(name))
        xs = rng.integers(0, s.width, n).astype(np.int32); ys = rng.integers(0, s.height, n).astype(np.int32)
        st = rng.integers(0, 2 ** 40, n).astype(np.uint64)
        rec = np.array([s.shift_sample(int(x), int(y), int(t)) for x, y, t in zip(xs, ys, st)])
        hit = (rec[:, 3:6] > 0).any(axis=1)
        shifted = (rec[:, 6:18].reshape(n, 4, 3) > 0).any(axis=2)
        print("%s: %d samples, base paths on a light %.3f, successful shifts among them %.3f, finite %s" % (
            name, n, hit.mean(), shifted[hit].mean() if hit.any() else 0.0, np.isfinite(rec).all()))
        np.savez_compressed(os.path.join(HERE, "shift_%s.npz" % name), xs=xs, ys=ys, streams=st, rec=rec)


if __name__ == "__main__":
    main()
