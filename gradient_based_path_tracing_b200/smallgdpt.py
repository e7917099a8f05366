"""The reference's stand-alone program small_gdpt.cpp on the GPU.

`render()` is /root/reference/small_gdpt.cpp's main() (:439-563): its built-in scene of eight spheres, `spp` base paths per
pixel with four shift-mapped offset paths each, the merged gradients and the screened-Poisson reconstruction.  The images
come back as the program holds them before it writes its four PFM files (:543-561): float64, [height][width][3], top row
first, dx / dy signed (`pfm_images()` gives what the files hold: float32, gradients as absolute values).
Everything runs in libgdpt.so (csrc/gdpt_smallgdpt_f64.cu); there is no CPU path here.
"""
import ctypes as C

import numpy as np

from . import _capi

ROWS, PIXELS = 0, 1
_MODES = {"rows": ROWS, "pixels": PIXELS, ROWS: ROWS, PIXELS: PIXELS}


def render(width=1024, height=768, spp=100, rng="rows", device=0):
    """width / height default to the program's film (small_gdpt.cpp:55), spp to its default sample count (:442).
    rng="rows": the program's erand48 streams, one per film row (:462); rng="pixels": one stream and one thread per pixel."""
    if rng not in _MODES:
        raise ValueError("rng must be 'rows' or 'pixels'")
    L = _capi.lib()
    w, h = int(width), int(height)
    ms = C.c_double(0.0)
    if not (2 <= w <= 16384 and 2 <= h <= 16384):        # no images to allocate: let the library state its limits
        _capi.check(L.gdpt_smallgdpt_render(int(device), w, h, int(spp), _MODES[rng], None, C.byref(ms)))
        raise ValueError("film size out of range")
    imgs = [_capi.pinned_empty((h, w, 3)) for _ in range(4)]
    out = _capi.GdptImages(*[a.ctypes.data_as(_capi.dp) for a in imgs])
    _capi.check(L.gdpt_smallgdpt_render(int(device), w, h, int(spp), _MODES[rng], C.byref(out), C.byref(ms)))
    return {"primal": imgs[0], "dx": imgs[1], "dy": imgs[2], "poisson": imgs[3], "render_ms": ms.value}


def pfm_images(r):
    """The four images as small_gdpt.cpp writes them (:543-561): image, image_dx, image_dy, image_poisson."""
    return {"image": r["primal"].astype(np.float32), "image_dx": np.abs(r["dx"]).astype(np.float32),
            "image_dy": np.abs(r["dy"]).astype(np.float32), "image_poisson": r["poisson"].astype(np.float32)}
