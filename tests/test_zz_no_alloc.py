"""gdpt_poisson_device_noalloc (include/gdpt.h): the screened-Poisson solve with its scratch in a fixed __device__ array, for
callers that cannot call cudaMalloc.  Called through the C ABI on device buffers and compared with the existing entry point
gdpt_poisson (same kernels, allocated work buffers) within the Poisson tolerance of this suite (1e-9 relative L2; gate 1e-5),
on calls in a row with different sizes so stale scratch contents would show.  Collected last (file name) on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def rel_l2(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def _noalloc(g, torch, u, gx, gy, alpha=0.04, stream=None):
    lib = g._capi.lib()
    h, w, _ = u.shape
    dev = torch.device("cuda", 0)
    t = [torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64).reshape(-1)).to(dev) for a in (u, gx, gy)]
    out = torch.full((h * w * 3,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    vp = C.c_void_p
    rc = lib.gdpt_poisson_device_noalloc(0, w, h, vp(t[0].data_ptr()), vp(t[1].data_ptr()), vp(t[2].data_ptr()), alpha, vp(out.data_ptr()),
                                         vp(stream.cuda_stream if stream is not None else None))
    g._capi.check(rc)
    if stream is not None:
        stream.synchronize()
    torch.cuda.synchronize()
    return out.cpu().numpy().reshape(h, w, 3)


def test_noalloc_poisson_equals_the_allocating_entry_point(g):
    import torch
    rng = np.random.default_rng(77)
    side = torch.cuda.Stream()
    # two calls in a row (then two more): a large image first, then smaller ones that reuse the same scratch
    for i, (h, w) in enumerate([(256, 256), (48, 64), (131, 77), (2, 2)]):
        u = rng.random((h, w, 3)); gx = rng.normal(0, 0.1, (h, w, 3)); gy = rng.normal(0, 0.1, (h, w, 3))
        want = g.fourier_solve(u, gx, gy, 0.04)
        got = _noalloc(g, torch, u, gx, gy, 0.04, stream=side if i % 2 else None)
        assert np.isfinite(got).all()
        assert rel_l2(got, want) < 1e-9, (h, w)
    # the same inputs twice in a row give the same bits
    a = _noalloc(g, torch, u, gx, gy)
    b = _noalloc(g, torch, u, gx, gy)
    assert np.array_equal(a, b)


def test_noalloc_poisson_rejects_sizes_beyond_the_fixed_scratch(g):
    import torch
    lib = g._capi.lib()
    z = torch.zeros(8, dtype=torch.float64, device="cuda")
    p = C.c_void_p(z.data_ptr())
    for w, h in ((257, 16), (16, 257), (1920, 1080)):
        rc = lib.gdpt_poisson_device_noalloc(0, w, h, p, p, p, 0.04, p, None)
        assert rc == -1
        msg = lib.gdpt_last_error().decode()
        assert "exceeds the fixed scratch" in msg and "256" in msg
    assert lib.gdpt_poisson_device_noalloc(0, 1, 8, p, p, p, 0.04, p, None) == -1
