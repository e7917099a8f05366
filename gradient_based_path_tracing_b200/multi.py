"""Multi-GPU gradpath: one process per GPU, image split into contiguous row bands with the
scene replicated (the thread-pool tile loop of /root/reference/src/render.cpp:271-331 becomes
per-rank bands), then ONE exchange: the five accumulation images are summed onto rank 0
(bands are disjoint, every rank holds zeros elsewhere, so the sum IS the gather and it also
delivers the cx1/cy1 halo rows the merge at render.cpp:345-349 needs).  Rank 0 merges and
runs the screened-Poisson solve.  torch.distributed (NCCL on GPUs, gloo in the CPU tests) is
plumbing only; all arithmetic stays in libgdpt.so.

Large films (BASELINE configs[4], 3840x2160) use the DISTRIBUTED solve instead: every rank keeps its
band, exchanges one halo row with each neighbour, and the 2-D DCT-I of fourierSolve
(/root/reference/src/render.cpp:172-254) runs as row transforms on row bands, an all-to-all
transpose, column transforms + spectral divide on column bands, an all-to-all back, and row
transforms again (`DistributedPoisson`; the local stages are gdpt_poisson_dist_* of include/gdpt.h).
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


def column_band_width(width, n_ranks):
    """Width of the column bands of the distributed solve: equal bands, the last takes the remainder."""
    bw = -(-width // n_ranks)
    if bw * (n_ranks - 1) >= width:
        bw = width // n_ranks
    if bw < 1:
        raise ValueError("image narrower than the number of ranks")
    return bw


def column_bands(width, n_ranks):
    bw = column_band_width(width, n_ranks)
    return [(s * bw, width if s == n_ranks - 1 else (s + 1) * bw) for s in range(n_ranks)]


class CudaStages:
    """The local stages of the distributed solve on this rank's GPU (libgdpt.so through the C ABI).
    Buffers are torch CUDA tensors of float64."""

    def __init__(self, ctx, capi):
        self.ctx, self.capi = ctx, capi

    @staticmethod
    def _p(t):
        return C.c_void_p(t.data_ptr())

    def _geom(self, g):
        return self.capi.GdptDistGeom(g["w"], g["h"], g["n_ranks"], g["rank"], g["y0"], g["y1"], g["band_w"])

    def merge_rows(self, g, ya, yb, cx0, cx1, cy0, cy1, cx, cy):
        lib, p = self.capi.lib(), self._p
        self.capi.check(lib.gdpt_merge_rows_device(self.ctx, g["w"], g["h"], ya, yb, p(cx0), p(cx1), p(cy0), p(cy1), p(cx), p(cy)))

    def rows_forward(self, g, u, gx, gy, alpha, send, dc_partial):
        lib, p = self.capi.lib(), self._p
        self.capi.check(lib.gdpt_poisson_dist_rows_forward(self.ctx, C.byref(self._geom(g)), p(u), p(gx), p(gy), alpha, p(send), p(dc_partial)))

    def columns(self, g, alpha, dc, recv, send_back):
        lib, p = self.capi.lib(), self._p
        self.capi.check(lib.gdpt_poisson_dist_columns(self.ctx, C.byref(self._geom(g)), alpha, p(dc), p(recv), p(send_back)))

    def rows_inverse(self, g, recv_back, out):
        lib, p = self.capi.lib(), self._p
        self.capi.check(lib.gdpt_poisson_dist_rows_inverse(self.ctx, C.byref(self._geom(g)), p(recv_back), p(out)))

    def sync(self):
        self.capi.check(self.capi.lib().gdpt_synchronize(self.ctx))


class DistributedPoisson:
    """merge (render.cpp:340-350) + fourierSolve (render.cpp:172-254) with the image split over the ranks.

    Rank r holds rows `bands[r]` of the five accumulation images inside full-size buffers.  Collectives:
    one halo row each way between neighbours, a 3-double all-reduce (the DC sums, render.cpp:199-211), two
    all-to-all transposes, and (optionally) the gather of the finished bands onto rank 0.  `stages`
    supplies the local arithmetic (CudaStages in the product; the CPU tests plug a numpy restatement in to
    exercise this plumbing over gloo).  `sync` is called around collectives so that the stage kernels (on
    the context's stream) and the collectives (on torch's) are ordered."""

    def __init__(self, width, height, bands, dist, torch, stages, device=None):
        self.w, self.h, self.bands, self.dist, self.torch, self.stages = width, height, list(bands), dist, torch, stages
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        if len(self.bands) != self.world or any(b[1] <= b[0] for b in self.bands):
            raise ValueError("the distributed solve needs one non-empty row band per rank")
        self.cols = column_bands(width, self.world)
        self.band_w = column_band_width(width, self.world)
        y0, y1 = self.bands[self.rank]
        x0, x1 = self.cols[self.rank]
        self.geom = dict(w=width, h=height, n_ranks=self.world, rank=self.rank, y0=y0, y1=y1, band_w=self.band_w)
        kw = dict(dtype=torch.float64, device=device)
        rows, wme = y1 - y0, x1 - x0
        self.send = torch.zeros(3 * rows * width, **kw)          # row band, chunked by destination column band
        self.recv = torch.zeros(3 * height * wme, **kw)          # column band [h][3*w_me]
        self.send_back = torch.zeros(3 * height * wme, **kw)
        self.recv_back = torch.zeros(3 * rows * width, **kw)
        self.dc = torch.zeros(3, **kw)
        self.row_splits = [rows * 3 * (c1 - c0) for c0, c1 in self.cols]           # what this row band sends to / gets back from rank s
        self.col_splits = [(b1 - b0) * 3 * wme for b0, b1 in self.bands]           # what this column band gets from / returns to rank r

    def _sync(self):
        self.stages.sync()
        if self.torch.cuda.is_available() and self.send.is_cuda:
            self.torch.cuda.current_stream().synchronize()

    def halo(self, cy0, cy1):
        """cy1 row y0-1 from rank-1 and cy0 row y1 from rank+1, into the full-size buffers (render.cpp:348)."""
        dist, r, n, row = self.dist, self.rank, self.world, self.w * 3
        y0, y1 = self.bands[r]
        ops = []
        if r + 1 < n:
            ops.append(dist.P2POp(dist.isend, cy1[(y1 - 1) * row:y1 * row], r + 1))
            ops.append(dist.P2POp(dist.irecv, cy0[y1 * row:(y1 + 1) * row], r + 1))
        if r > 0:
            ops.append(dist.P2POp(dist.isend, cy0[y0 * row:(y0 + 1) * row], r - 1))
            ops.append(dist.P2POp(dist.irecv, cy1[(y0 - 1) * row:y0 * row], r - 1))
        if ops:
            for q in dist.batch_isend_irecv(ops):
                q.wait()

    def solve(self, img, cx0, cy0, cx1, cy1, cx, cy, out, alpha=0.04):
        """All buffers are full-size (W*H*3) float64; on return rows bands[rank] of cx, cy, out are final."""
        dist, st, g = self.dist, self.stages, self.geom
        y0, y1 = self.bands[self.rank]
        self._sync()
        if self.world > 1:
            self.halo(cy0, cy1)
            self._sync()          # the halo rows must have landed before the merge kernel (context stream) reads them
        st.merge_rows(g, y0, min(y1 + 1, self.h), cx0, cx1, cy0, cy1, cx, cy)
        st.rows_forward(g, img, cx, cy, alpha, self.send, self.dc)
        self._sync()
        if self.world > 1:
            dist.all_reduce(self.dc, op=dist.ReduceOp.SUM)
            dist.all_to_all_single(self.recv, self.send, output_split_sizes=self.col_splits, input_split_sizes=self.row_splits)
            self._sync()
            st.columns(g, alpha, self.dc, self.recv, self.send_back)
            self._sync()
            dist.all_to_all_single(self.recv_back, self.send_back, output_split_sizes=self.row_splits, input_split_sizes=self.col_splits)
            self._sync()
            st.rows_inverse(g, self.recv_back, out)
        else:
            st.columns(g, alpha, self.dc, self.send, self.send_back)
            st.rows_inverse(g, self.send_back, out)
        self._sync()

    def gather_rows(self, buffers, dst=0):
        """Bring the row bands of full-size images together on rank `dst` (contiguous row slices, point to point)."""
        dist, row = self.dist, self.w * 3
        if self.world == 1:
            return
        ops = []
        for b in buffers:
            if self.rank == dst:
                for r, (b0, b1) in enumerate(self.bands):
                    if r != dst:
                        ops.append(dist.P2POp(dist.irecv, b[b0 * row:b1 * row], r))
            else:
                b0, b1 = self.bands[self.rank]
                ops.append(dist.P2POp(dist.isend, b[b0 * row:b1 * row], dst))
        for q in dist.batch_isend_irecv(ops):
            q.wait()


def share_bvh_from_rank0(scene, device, dist, torch):
    """Rank 0's BVH (gdpt_bvh_info) -> every other rank's context (gdpt_bvh_adopt): sizes in one broadcast, the 128-byte
    nodes and the 48-byte primitive records in two more, device to device."""
    dev = torch.device("cuda", device)
    rank = dist.get_rank()
    meta = torch.zeros(3, dtype=torch.int64, device=dev)
    if rank == 0:
        nn, npr, depth, pn, pp = scene.bvh_info(device)
        meta[0], meta[1], meta[2] = nn, npr, depth
    dist.broadcast(meta, src=0)
    nn, npr, depth = (int(v) for v in meta.tolist())
    nodes = torch.empty(max(nn, 1) * 128, dtype=torch.uint8, device=dev)
    prims = torch.empty(max(npr, 1) * 48, dtype=torch.uint8, device=dev)
    if rank == 0:
        nodes.copy_(_as_u8(torch, pn, nn * 128, dev)); prims.copy_(_as_u8(torch, pp, npr * 48, dev))
    dist.broadcast(nodes, src=0)
    dist.broadcast(prims, src=0)
    torch.cuda.current_stream().synchronize()
    if rank != 0:
        scene.adopt_bvh(nodes.data_ptr(), nn, prims.data_ptr(), npr, depth, device=device)


def _as_u8(torch, ptr, nbytes, dev):
    """A torch view of `nbytes` of device memory at `ptr` (memory owned by the gdpt context)."""
    class _Cai:
        __cuda_array_interface__ = {"shape": (max(nbytes, 1),), "typestr": "|u1", "data": (ptr, False), "version": 2}
    return torch.as_tensor(_Cai(), device=dev)


class DistributedRenderer:
    """render() across the ranks of an initialised torch.distributed process group."""

    def __init__(self, scene, device, dist, torch, solver="auto", share_bvh="auto", bvh=None):
        """solver: "gather" (sum the bands onto rank 0 and solve there), "distributed" (all-to-all transposed
        solve on all ranks) or "auto" (distributed from 3840x2160 pixels up when every rank has rows).
        share_bvh: rank 0 builds the BVH, the others create their context without one and receive rank 0's through a
        broadcast (the ranks of a box share its host cores: N identical builds side by side cost N times the core time).
        "auto": from 4 ranks up — measured on B200, end-to-end step at 64 spp: 8 ranks 320 -> 309 ms, 2 ranks 379 -> 387 ms
        (two builds do not contend on 16 cores, and the other rank waits for rank 0's)."""
        from . import _capi
        self.scene, self.device, self.dist, self.torch, self._capi = scene, device, dist, torch, _capi
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.bands = partition_rows(scene.height, self.world)
        # which builder: a rank's share of the planned work decides like gdpt_create does for the whole film (include/gdpt.h:
        # the host's better tree pays from ~1e8 pixel-samples).  Device builds are per rank (9 ms, no host cores involved);
        # a host build is done once on rank 0 and broadcast from 4 ranks up (N builds side by side share the box's cores).
        share_of_work = scene.width * scene.height * max(1, scene.options.samples_per_pixel) / self.world
        self.bvh_builder = bvh or ("device" if share_of_work < 1e8 else "host")      # bvh: "host" | "device" forces one
        if share_bvh == "auto":
            share_bvh = self.world >= 4
        if device in scene._ctx:
            self.ctx = scene.context(device)
        elif self.bvh_builder == "device":
            self.ctx = scene.context(device, bvh="device")
        elif share_bvh and self.world > 1:
            self.ctx = scene.context(device, build_bvh=self.rank == 0, bvh="host")
            share_bvh_from_rank0(scene, device, dist, torch)
        else:
            self.ctx = scene.context(device, bvh="host")
        n = scene.height * scene.width * 3
        dev = torch.device("cuda", device)
        self.acc = [torch.zeros(n, dtype=torch.float64, device=dev) for _ in range(5)]
        self.cx = torch.zeros(n, dtype=torch.float64, device=dev)
        self.cy = torch.zeros(n, dtype=torch.float64, device=dev)
        self.out = torch.zeros(n, dtype=torch.float64, device=dev)
        if solver == "auto":
            solver = "distributed" if (self.world > 1 and scene.width * scene.height >= 3840 * 2160) else "gather"
        if solver == "distributed" and (self.world == 1 or any(b[1] <= b[0] for b in self.bands)):
            solver = "gather"
        self.solver = solver
        # gather solve: the film is shared as interleaved 4-row strips (strip i belongs to rank i mod N — what the reference's
        # dynamic tile queue achieves, parallel.cpp:183-237): contiguous bands of Sponza differ by up to 8 % in cost (sky
        # openings and arches against the floor), strips by well under 1 %.  The distributed solve needs contiguous row bands.
        self.interleave = solver == "gather" and self.world > 1
        self.last_ms = {}
        self.dsolve = None
        if solver == "distributed":
            self.dsolve = DistributedPoisson(scene.width, scene.height, self.bands, dist, torch, CudaStages(self.ctx, _capi), device=dev)

    def render_path(self, spp, precision="f32", seed=0, to_host=False):
        """path_render (render.cpp:74-117) over the ranks: each renders its row band, the bands are summed onto rank 0."""
        from .render import _params
        lib, check, torch = self._capi.lib(), self._capi.check, self.torch
        y0, y1 = self.bands[self.rank]
        img = self.acc[0]
        img.zero_()
        torch.cuda.current_stream().synchronize()
        if y1 > y0:
            p = _params(self.scene, spp=spp, precision=precision, seed=seed, rows=(y0, y1))
            if y0 == 0:
                p.y0 = 0
            check(lib.gdpt_render_path_accumulate_device(self.ctx, C.byref(p), C.c_void_p(img.data_ptr())))
            check(lib.gdpt_synchronize(self.ctx))
        if self.world > 1:
            reduce_bands([img], self.dist, dst=0)
            torch.cuda.current_stream().synchronize()
        if self.rank == 0 and to_host:
            return img.cpu().numpy().reshape(self.scene.height, self.scene.width, 3)
        return None

    def _to_host(self, named):
        """Device images -> page-locked host arrays from the binding's pinned pool (_capi.pinned_empty: the blocks go back to
        the pool when the arrays are collected), so the copy runs at PCIe speed."""
        torch = self.torch
        shape = (self.scene.height, self.scene.width, 3)
        out = {}
        for k, v in named:
            a = self._capi.pinned_empty(shape)
            torch.from_numpy(a.reshape(-1)).copy_(v, non_blocking=True)
            out[k] = a
        torch.cuda.current_stream().synchronize()
        return out

    def render(self, spp, precision="f32", seed=0, data_cost=0.04, to_host=False, offsets="reference"):
        """Times of the last call (CUDA events on this rank's stream, ms) are left in self.last_ms:
        render (this rank's share of the samples), exchange (reduce / halo + transposes + gather), solve, copy."""
        from .render import _params
        lib, check = self._capi.lib(), self._capi.check
        torch = self.torch
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        y0, y1 = self.bands[self.rank]
        for b in self.acc:
            b.zero_()
        torch.cuda.current_stream().synchronize()
        ev[0].record()
        if self.interleave:
            p = _params(self.scene, spp=spp, precision=precision, seed=seed, offsets=offsets, strips=(self.world, self.rank))
        elif y1 > y0:
            p = _params(self.scene, spp=spp, precision=precision, seed=seed, rows=(y0, y1), offsets=offsets)
            if y0 == 0:
                p.y0 = 0
        else:
            p = None
        if p is not None:
            ptrs = [C.c_void_p(b.data_ptr()) for b in self.acc]
            check(lib.gdpt_render_accumulate_device(self.ctx, C.byref(p), *ptrs))
            check(lib.gdpt_synchronize(self.ctx))
        ev[1].record()
        res = None
        if self.dsolve is not None:
            a = self.acc
            self.dsolve.solve(a[0], a[1], a[2], a[3], a[4], self.cx, self.cy, self.out, alpha=data_cost)
            self.dsolve.gather_rows([a[0], self.cx, self.cy, self.out] if to_host else [self.out], dst=0)
            torch.cuda.current_stream().synchronize()
            ev[3].record()
            if self.rank == 0 and to_host:
                res = self._to_host((("primal", a[0]), ("dx", self.cx), ("dy", self.cy), ("poisson", self.out)))
            ev[4].record(); ev[4].synchronize()
            self.last_ms = {"render": ev[0].elapsed_time(ev[1]), "exchange+solve": ev[1].elapsed_time(ev[3]), "copy": ev[3].elapsed_time(ev[4])}
            return res
        if self.world > 1:
            reduce_bands(self.acc, self.dist, dst=0)
            torch.cuda.current_stream().synchronize()
        ev[2].record()
        if self.rank == 0:
            w, h = self.scene.width, self.scene.height
            a = self.acc
            check(lib.gdpt_merge_device(self.ctx, w, h, C.c_void_p(a[1].data_ptr()), C.c_void_p(a[3].data_ptr()),
                                        C.c_void_p(a[2].data_ptr()), C.c_void_p(a[4].data_ptr()),
                                        C.c_void_p(self.cx.data_ptr()), C.c_void_p(self.cy.data_ptr())))
            check(lib.gdpt_poisson_device(self.ctx, w, h, C.c_void_p(a[0].data_ptr()), C.c_void_p(self.cx.data_ptr()),
                                          C.c_void_p(self.cy.data_ptr()), data_cost, C.c_void_p(self.out.data_ptr())))
            check(lib.gdpt_synchronize(self.ctx))
            ev[3].record()
            if to_host:
                res = self._to_host((("primal", a[0]), ("dx", self.cx), ("dy", self.cy), ("poisson", self.out)))
        else:
            ev[3].record()
        ev[4].record(); ev[4].synchronize()
        self.last_ms = {"render": ev[0].elapsed_time(ev[1]), "exchange": ev[1].elapsed_time(ev[2]), "solve": ev[2].elapsed_time(ev[3]),
                        "copy": ev[3].elapsed_time(ev[4])}
        return res
