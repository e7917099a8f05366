"""B200-native gradient-domain path tracer: drop-in for lajolla's `gradpath` integrator.

The product is libgdpt.so (CUDA, sm_100a) behind include/gdpt.h; this package is the thin
host-side mirror of the reference's `render(const Scene&)` interface for that path.
"""
from .scene import Integrator, RenderOptions, Scene  # noqa: F401
from .render import (render, gradient_path_render, gradient_path_render_multi, path_render, aux_render, path_samples, sample_light, occluded, fourier_solve, reconstruct_l1, grad_samples, intersect, sample_primary,  # noqa: F401
                     bsdf, pcg_fill, scene_bounds, stats)
