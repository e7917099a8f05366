"""Multi-GPU gradpath: one process per GPU, image split into contiguous row bands with the
scene replicated (the thread-pool tile loop of /root/reference/src/render.cpp:271-331 becomes
per-rank bands), then ONE exchange: the five accumulation images are summed onto rank 0
(bands are disjoint, every rank holds zeros elsewhere, so the sum IS the gather and it also
delivers the cx1/cy1 halo rows the merge at render.cpp:345-349 needs).  Rank 0 merges and
runs the screened-Poisson solve.  torch.distributed (NCCL on GPUs, gloo in the CPU tests) is
plumbing only; all arithmetic stays in libgdpt.so.
"""
import ctypes as C

import numpy as np


def partition_rows(height, n_ranks):
    """Contiguous row bands [y0, y1) per rank; band starts are multiples of 4 (the wavefront's
    8x4 pixel tiles), sizes differ by at most one 4-row strip."""
    strips = (height + 3) // 4
    base, extra = divmod(strips, n_ranks)
    bands, s = [], 0
    for r in range(n_ranks):
        k = base + (1 if r < extra else 0)
        y0, y1 = min(4 * s, height), min(4 * (s + k), height)
        bands.append((y0, y1))
        s += k
    bands[-1] = (bands[-1][0], height)
    return bands


def reduce_bands(buffers, dist, dst=0):
    """Sum the per-rank full-size images onto rank `dst` (the path's only collective)."""
    for b in buffers:
        dist.reduce(b, dst=dst, op=dist.ReduceOp.SUM)


class DistributedRenderer:
    """render() across the ranks of an initialised torch.distributed process group."""

    def __init__(self, scene, device, dist, torch):
        from . import _capi
        self.scene, self.device, self.dist, self.torch, self._capi = scene, device, dist, torch, _capi
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.bands = partition_rows(scene.height, self.world)
        self.ctx = scene.context(device)
        n = scene.height * scene.width * 3
        dev = torch.device("cuda", device)
        self.acc = [torch.zeros(n, dtype=torch.float64, device=dev) for _ in range(5)]
        self.cx = torch.zeros(n, dtype=torch.float64, device=dev)
        self.cy = torch.zeros(n, dtype=torch.float64, device=dev)
        self.out = torch.zeros(n, dtype=torch.float64, device=dev)

    def render(self, spp, precision="f32", seed=0, data_cost=0.04, to_host=False):
        from .render import _params
        lib, check = self._capi.lib(), self._capi.check
        torch = self.torch
        y0, y1 = self.bands[self.rank]
        for b in self.acc:
            b.zero_()
        torch.cuda.current_stream().synchronize()
        if y1 > y0:
            p = _params(self.scene, spp=spp, precision=precision, seed=seed, rows=(y0, y1))
            if y0 == 0:
                p.y0 = 0
            ptrs = [C.c_void_p(b.data_ptr()) for b in self.acc]
            check(lib.gdpt_render_accumulate_device(self.ctx, C.byref(p), *ptrs))
            check(lib.gdpt_synchronize(self.ctx))
        if self.world > 1:
            reduce_bands(self.acc, self.dist, dst=0)
            torch.cuda.current_stream().synchronize()
        res = None
        if self.rank == 0:
            w, h = self.scene.width, self.scene.height
            a = self.acc
            check(lib.gdpt_merge_device(self.ctx, w, h, C.c_void_p(a[1].data_ptr()), C.c_void_p(a[3].data_ptr()),
                                        C.c_void_p(a[2].data_ptr()), C.c_void_p(a[4].data_ptr()),
                                        C.c_void_p(self.cx.data_ptr()), C.c_void_p(self.cy.data_ptr())))
            check(lib.gdpt_poisson_device(self.ctx, w, h, C.c_void_p(a[0].data_ptr()), C.c_void_p(self.cx.data_ptr()),
                                          C.c_void_p(self.cy.data_ptr()), data_cost, C.c_void_p(self.out.data_ptr())))
            check(lib.gdpt_synchronize(self.ctx))
            if to_host:
                shape = (h, w, 3)
                res = {k: v.cpu().numpy().reshape(shape) for k, v in
                       (("primal", a[0]), ("dx", self.cx), ("dy", self.cy), ("poisson", self.out))}
        return res
