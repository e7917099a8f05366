"""Host-side mirror of the reference's operator interface for the hot path:

    Image3 render(const Scene &scene)          /root/reference/src/render.h:9
    -> Integrator::GradPath arm                 /root/reference/src/render.cpp:387-388

`render(scene)` returns the Poisson-reconstructed image like the reference does;
`gradient_path_render(scene)` additionally returns the primal / dx / dy buffers the
reference computes and then drops (render.cpp:355-369).  All arithmetic runs in
libgdpt.so on the GPU; an integrator other than GradPath raises like the reference's
`assert(false)` arm (render.cpp:390-393).
"""
import ctypes as C

import numpy as np

from . import _capi
from .scene import Integrator, Scene


class RenderResult(dict):
    """primal, dx, dy, poisson: (H, W, 3) float64 arrays; stats: dict"""
    __getattr__ = dict.__getitem__


def _params(scene, spp=None, precision="f64", seed=0, sample0=0, max_depth=None, rr_depth=None, data_cost=0.04,
            rows=None, wave_samples=0, offsets="reference", strips=None):
    p = _capi.GdptRenderParams()
    p.spp = int(spp if spp is not None else scene.options.samples_per_pixel)
    p.sample0 = int(sample0)
    p.max_depth = -2 if max_depth is None else int(max_depth)
    p.rr_depth = -2 if rr_depth is None else int(rr_depth)
    p.seed = int(seed)
    p.data_cost = float(data_cost)
    p.precision = {"f64": _capi.PRECISION_F64, "f32": _capi.PRECISION_F32}[precision]
    p.y0, p.y1 = (0, 0) if rows is None else (int(rows[0]), int(rows[1]))
    p.wave_samples = int(wave_samples)
    p.offset_mode = {"reference": _capi.OFFSETS_REFERENCE, "shift": _capi.OFFSETS_SHIFT}[offsets]
    p.strip_count, p.strip_phase = (0, 0) if strips is None else (int(strips[0]), int(strips[1]))    # (count, phase): every count-th 4-row strip
    return p


def stats(scene, device=0):
    s = _capi.GdptStats()
    _capi.check(_capi.lib().gdpt_get_stats(scene.context(device), C.byref(s)))
    return {k: getattr(s, k) for k, _ in s._fields_}


def gradient_path_render(scene, device=0, outputs=("primal", "dx", "dy", "poisson"), pinned=True, **kw):
    """gradient_path_render (render.cpp:257-370) on the GPU; returns all four buffers (page-locked host arrays by default)."""
    ctx = scene.context(device)
    p = _params(scene, **kw)
    imgs = _capi.GdptImages()
    res = RenderResult()
    for name in outputs:
        a = _capi.pinned_empty((scene.height, scene.width, 3)) if pinned else np.empty((scene.height, scene.width, 3), dtype=np.float64)
        res[name] = a
        setattr(imgs, name, _capi.as_dp(a))
    _capi.check(_capi.lib().gdpt_render(ctx, C.byref(p), C.byref(imgs)))
    res["stats"] = stats(scene, device)
    return res


def gradient_path_render_multi(scene, n_gpus, outputs=("primal", "dx", "dy", "poisson"), devices=None, **kw):
    """gradient_path_render over `n_gpus` GPUs of this box inside this process (gdpt_multi_*: one host thread per device,
    interleaved 4-row strips, NVLink peer copies to the first device, merge + solve there)."""
    lib = _capi.lib()
    h = C.c_void_p()
    d = scene.desc()
    dev = (C.c_int32 * n_gpus)(*devices) if devices is not None else None
    _capi.check(lib.gdpt_multi_create(C.byref(d), n_gpus, dev, C.byref(h)))
    try:
        p = _params(scene, **kw)
        imgs = _capi.GdptImages()
        res = RenderResult()
        for name in outputs:
            a = _capi.pinned_empty((scene.height, scene.width, 3))
            res[name] = a
            setattr(imgs, name, _capi.as_dp(a))
        _capi.check(lib.gdpt_multi_render(h, C.byref(p), C.byref(imgs)))
        s = _capi.GdptStats()
        _capi.check(lib.gdpt_multi_get_stats(h, C.byref(s)))
        res["stats"] = {k: getattr(s, k) for k, _ in s._fields_}
    finally:
        lib.gdpt_multi_destroy(h)
    return res


def path_render(scene, device=0, **kw):
    """path_render (render.cpp:74-117): the plain NEE + MIS path tracer, the metric's equal-time comparator.
    Returns the (H, W, 3) image; scenes with an environment-map light raise GdptError (not built)."""
    ctx = scene.context(device)
    p = _params(scene, **kw)
    img = np.zeros((scene.height, scene.width, 3), dtype=np.float64)
    _capi.check(_capi.lib().gdpt_render_path(ctx, C.byref(p), _capi.as_dp(img)))
    return img


AUX_INTEGRATORS = (Integrator.Depth, Integrator.ShadingNormal, Integrator.MeanCurvature, Integrator.RayDifferential, Integrator.MipmapLevel)


def aux_render(scene, integrator=None, precision="f64", device=0):
    """aux_render (render.cpp:15-72): depth / shading normal / mean curvature / ray differential / mip level of the
    primary vertex of every pixel centre, (H, W, 3)."""
    kind = Integrator(scene.options.integrator if integrator is None else integrator)
    if kind not in AUX_INTEGRATORS:
        raise ValueError("%s is not an auxiliary integrator" % kind.name)
    img = np.zeros((scene.height, scene.width, 3), dtype=np.float64)
    _capi.check(_capi.lib().gdpt_render_aux(scene.context(device), int(kind), _prec(precision), _capi.as_dp(img)))
    return img


def render(scene, device=0, **kw):
    """Image3 render(const Scene&) (render.cpp:373-395): dispatch on scene.options.integrator; GradPath and Path
    and the
    auxiliary buffers live here; VolPath ends like the reference's `assert(false)` arm."""
    if scene.options.integrator in AUX_INTEGRATORS:
        return aux_render(scene, precision=kw.get("precision", "f64"), device=device)
    if scene.options.integrator == Integrator.Path:
        return path_render(scene, device, **kw)
    if scene.options.integrator != Integrator.GradPath:
        raise AssertionError("render(): integrator %s is outside the gradpath hot path (render.cpp:390-393)" % scene.options.integrator.name)
    return gradient_path_render(scene, device, outputs=("poisson",), **kw)["poisson"]


def fourier_solve(u, gx, gy, data_cost=0.04, device=0):
    """fourierSolve(width, height, imgData, imgGradX, imgGradY, dataCost, imgOut) (render.cpp:172-254)."""
    u, gx, gy = (_capi.f64(a) for a in (u, gx, gy))
    if u.ndim != 3 or u.shape[2] != 3 or gx.shape != u.shape or gy.shape != u.shape:
        raise ValueError("expected three (H, W, 3) arrays")
    h, w, _ = u.shape
    out = np.zeros_like(u)
    _capi.check(_capi.lib().gdpt_poisson(device, w, h, _capi.as_dp(u), _capi.as_dp(gx), _capi.as_dp(gy), data_cost, _capi.as_dp(out)))
    return out


def reconstruct_l1(u, gx, gy, data_cost=0.04, irls_iterations=10, cg_iterations=50, epsilon=0.05, cg_tolerance=1e-6, start="fourier", reweight_primal=False, device=0, return_info=False):
    """L1 reconstruction by iteratively reweighted least squares (gdpt_poisson_l1; no counterpart in the reference's code, whose
    only reconstruction is fourierSolve).  start: "fourier" (fourierSolve's image), "primal" (u) or an (H, W, 3) array.
    irls_iterations=0 solves the L2 problem (all weights 1) by conjugate gradients.  reweight_primal=False keeps the primal term
    quadratic (the image keeps the primal's mean); True is the all-L1 energy."""
    u, gx, gy = (_capi.f64(a) for a in (u, gx, gy))
    if u.ndim != 3 or u.shape[2] != 3 or gx.shape != u.shape or gy.shape != u.shape:
        raise ValueError("expected three (H, W, 3) arrays")
    h, w, _ = u.shape
    out = np.zeros_like(u)
    if isinstance(start, str):
        if start not in ("fourier", "primal"):
            raise ValueError("start must be 'fourier', 'primal' or an image")
        mode = _capi.L1_START_FOURIER if start == "fourier" else _capi.L1_START_PRIMAL
    else:
        s0 = _capi.f64(start)
        if s0.shape != u.shape:
            raise ValueError("the starting image must have the shape of u")
        out[...] = s0
        mode = _capi.L1_START_OUT
    prm = _capi.GdptL1Params(int(irls_iterations), int(cg_iterations), mode, 1 if reweight_primal else 0, float(data_cost), float(epsilon), float(cg_tolerance))
    info = np.zeros(4)
    _capi.check(_capi.lib().gdpt_poisson_l1(device, w, h, _capi.as_dp(u), _capi.as_dp(gx), _capi.as_dp(gy), C.byref(prm), _capi.as_dp(out), _capi.as_dp(info)))
    if return_info:
        return out, {"relative_residual": float(info[0]), "cg_steps": int(info[1]), "ms": float(info[2]), "eps": float(info[3])}
    return out


# ---- validation entry points (same device functions as the renderer) ----
def _prec(precision):
    return {"f64": _capi.PRECISION_F64, "f32": _capi.PRECISION_F32}[precision]


def sample_primary(scene, screen_pos, precision="f64", device=0):
    sp = _capi.f64(screen_pos).reshape(-1, 2)
    rays = np.zeros((sp.shape[0], 6))
    _capi.check(_capi.lib().gdpt_sample_primary(scene.context(device), _prec(precision), sp.shape[0], _capi.as_dp(sp), _capi.as_dp(rays)))
    return rays


def intersect(scene, rays, rdiff=None, precision="f64", device=0):
    rays = _capi.f64(rays).reshape(-1, 8)
    n = rays.shape[0]
    ids = np.zeros((n, 3), dtype=np.int32)
    vtx = np.zeros((n, 24))
    rd = _capi.f64(rdiff).reshape(-1, 2) if rdiff is not None else None
    _capi.check(_capi.lib().gdpt_intersect(scene.context(device), _prec(precision), n, _capi.as_dp(rays), _capi.as_dp(rd),
                                           ids.ctypes.data_as(_capi.ip), _capi.as_dp(vtx)))
    return ids, vtx


def bsdf(scene, mat_id, vtx, dirs, rnd, precision="f64", device=0):
    vtx, dirs, rnd = _capi.f64(vtx).reshape(-1, 15), _capi.f64(dirs).reshape(-1, 6), _capi.f64(rnd).reshape(-1, 3)
    out = np.zeros((vtx.shape[0], 10))
    _capi.check(_capi.lib().gdpt_bsdf(scene.context(device), _prec(precision), mat_id, vtx.shape[0], _capi.as_dp(vtx), _capi.as_dp(dirs),
                                      _capi.as_dp(rnd), _capi.as_dp(out)))
    return out


def grad_samples(scene, xs, ys, streams, device=0, **kw):
    """grad_path_tracing for explicit (x, y, rng stream) triples -> (n, 23) records."""
    xs = np.ascontiguousarray(xs, dtype=np.int32)
    ys = np.ascontiguousarray(ys, dtype=np.int32)
    st = np.ascontiguousarray(streams, dtype=np.uint64)
    rec = np.zeros((xs.size, 23))
    p = _params(scene, **kw)
    _capi.check(_capi.lib().gdpt_grad_samples(scene.context(device), C.byref(p), xs.size, xs.ctypes.data_as(_capi.ip),
                                              ys.ctypes.data_as(_capi.ip), st.ctypes.data_as(C.POINTER(C.c_uint64)), _capi.as_dp(rec)))
    return rec


def path_samples(scene, xs, ys, streams, device=0, **kw):
    """path_tracing (path_tracing.h:13-348) for explicit (x, y, rng stream) triples -> (n, 3) radiance."""
    xs = np.ascontiguousarray(xs, dtype=np.int32)
    ys = np.ascontiguousarray(ys, dtype=np.int32)
    st = np.ascontiguousarray(streams, dtype=np.uint64)
    rad = np.zeros((xs.size, 3))
    p = _params(scene, **kw)
    _capi.check(_capi.lib().gdpt_path_samples(scene.context(device), C.byref(p), xs.size, xs.ctypes.data_as(_capi.ip),
                                              ys.ctypes.data_as(_capi.ip), st.ctypes.data_as(C.POINTER(C.c_uint64)), _capi.as_dp(rad)))
    return rad


def sample_light(scene, ref_points, rnd, precision="f64", device=0):
    """sample_light + sample_point_on_light + pdf_point_on_light + light_pmf: rnd[n][4] = light_w, uv0, uv1, shape_w
    -> (n, 9) = light id, position3, normal3, pdf of the point, pmf of the light."""
    ref_points, rnd = _capi.f64(ref_points).reshape(-1, 3), _capi.f64(rnd).reshape(-1, 4)
    out = np.zeros((ref_points.shape[0], 9))
    _capi.check(_capi.lib().gdpt_sample_light(scene.context(device), _prec(precision), ref_points.shape[0], _capi.as_dp(ref_points),
                                              _capi.as_dp(rnd), _capi.as_dp(out)))
    return out


def occluded(scene, rays, precision="f64", device=0):
    """occluded (intersection.cpp:66-85): rays[n][8] = org, dir, tnear, tfar -> bool[n]."""
    rays = _capi.f64(rays).reshape(-1, 8)
    out = np.zeros(rays.shape[0], dtype=np.int32)
    _capi.check(_capi.lib().gdpt_occluded(scene.context(device), _prec(precision), rays.shape[0], _capi.as_dp(rays), out.ctypes.data_as(_capi.ip)))
    return out.astype(bool)


def pcg_fill(stream, n, seed=_capi.DEFAULT_SEED, device=0):
    out = np.zeros(n)
    raw = np.zeros(n, dtype=np.uint32)
    _capi.check(_capi.lib().gdpt_pcg_fill(device, stream, seed, n, _capi.as_dp(out), raw.ctypes.data_as(C.POINTER(C.c_uint32))))
    return out, raw


def scene_bounds(scene, device=0):
    b = np.zeros(4)
    _capi.check(_capi.lib().gdpt_scene_bounds(scene.context(device), _capi.as_dp(b)))
    return b
