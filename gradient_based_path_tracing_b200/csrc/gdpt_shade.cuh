// Shading half of the wavefront (see gdpt_wavefront.cuh for the overview).
#pragma once
#include "gdpt_wavefront.cuh"

namespace gdpt {

// A finished sample -> the GraidentPTRadiance record (intersection.h:65-77) and its
// accumulation (render.cpp:311-327).  contribS/wS for disabled offsets are 0 / 1.
template <typename R>
GD_D void finish_path(const WaveParams &wp, int px, int py, int id, const V3<R> &rad, const double contrib[3], double prob,
                      const double cS[4][3], const double wS[4]) {
    if (wp.rec_out) {
        double *o = wp.rec_out + (size_t)id * 23;
        o[0] = rad.x; o[1] = rad.y; o[2] = rad.z;
        o[3] = contrib[0]; o[4] = contrib[1]; o[5] = contrib[2];
        // record order: contribX0, contribX1, contribY0, contribY1 ; offsets here are X0,Y0,X1,Y1
        const int slot[4] = {0, 2, 1, 3};
        for (int k = 0; k < 4; k++) for (int c = 0; c < 3; c++) o[6 + 3 * slot[k] + c] = cS[k][c];
        o[18] = prob; o[19] = wS[0]; o[20] = wS[1]; o[21] = wS[2]; o[22] = wS[3];
        return;
    }
    if (!(prob > 0.0)) return;
    size_t p = ((size_t)py * wp.width + px) * 3;
    double spp = (double)wp.spp_total;
    double inv_spp = 1.0 / spp;                      // Vector3 / Real multiplies by the reciprocal
    double rd[3] = {(double)rad.x, (double)rad.y, (double)rad.z};
    double *bufs[4] = {wp.cx0, wp.cy0, wp.cx1, wp.cy1};
    const double sign[4] = {1.0, 1.0, -1.0, -1.0};  // X0,Y0: contrib - contribS ; X1,Y1: contribS - contrib
    if (wp.res) {                                    // one contiguous 120-byte record per path; summed in sample order by k_wave_reduce
        double *o = wp.res + (size_t)id * 15;
        for (int c = 0; c < 3; c++) {
            o[c] = rd[c] * inv_spp;
            for (int k = 0; k < 4; k++) o[3 + 3 * k + c] = sign[k] * (contrib[c] - cS[k][c]) * (wS[k] / (prob * spp));
        }
        return;
    }
    for (int c = 0; c < 3; c++) {
        if (rd[c] != 0.0) atomicAdd(wp.img + p + c, rd[c] * inv_spp);
        for (int k = 0; k < 4; k++) {
            double v = sign[k] * (contrib[c] - cS[k][c]) * (wS[k] / (prob * spp));
            if (v != 0.0) atomicAdd(bufs[k] + p + c, v);
        }
    }
}

template <typename R>
GD_D void finish_with_offsets(const WaveParams &wp, const PathRec<R> &pr, const OffRec<R> *off, double jac_num /* p2 area, or <0: jacobians stay 1 */) {
    double cS[4][3], wS[4];
    for (int k = 0; k < 4; k++) {
        cS[k][0] = cS[k][1] = cS[k][2] = 0.0; wS[k] = 1.0;
        if (off && (off->enabled >> k & 1)) {
            double jac = 1.0;
            bool ok = true;
            if (jac_num >= 0.0) {
                if (off->sampled >> k & 1) jac = jac_num / (double)off->p2s[k];
                else ok = false;              // replayed sample failed or had pdf <= 0: contribS = 0, w = 1
            }
            if (ok) {
                cS[k][0] = (double)off->contrib[k].x * jac; cS[k][1] = (double)off->contrib[k].y * jac; cS[k][2] = (double)off->contrib[k].z * jac;
                wS[k] = pr.prob / (pr.prob + 1.0 * jac);
            }
        }
    }
    finish_path<R>(wp, pr.px, pr.py, pr.id, pr.rad, pr.contrib, pr.prob, cS, wS);
}

// The accumulation of render.cpp:311-327 without atomics: thread per pixel slot of the wave (the 8x4-tile enumeration of
// k_generate), the samples of the wave added one after the other — in sample order, onto the running image value — so the
// five images are bit-reproducible and carry the summation order of the reference's own per-pixel loop.  Waves are reduced in
// wave order (one stream), pixels are disjoint between threads, strips between GPUs.
static __global__ void __launch_bounds__(256) k_wave_reduce(WaveParams wp, const double *res, double *img, double *cx0, double *cy0, double *cx1, double *cy1) {
    const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (p >= wp.npix_pad) return;
    const int tile = (int)(p >> 5), lane = (int)(p & 31);
    const int x = (tile % wp.tiles_x) * 8 + (lane & 7), y = wave_row(wp, tile / wp.tiles_x, lane);
    if (x >= wp.width || y >= wp.y1) return;
    const size_t pix = ((size_t)y * wp.width + x) * 3;
    double *bufs[5] = {img, cx0, cy0, cx1, cy1};
    double acc[15];
#pragma unroll
    for (int k = 0; k < 15; k++) acc[k] = bufs[k / 3][pix + k % 3];
    for (int s = 0; s < wp.n_samples; s++) {
        const double *r = res + ((size_t)s * wp.npix_pad + p) * 15;
#pragma unroll
        for (int k = 0; k < 15; k++) acc[k] += r[k];
    }
#pragma unroll
    for (int k = 0; k < 15; k++) bufs[k / 3][pix + k % 3] = acc[k];
}

// Start of a loop iteration at `vtx` (path_tracing.h:528-564 + eval/pdf of :762-763, which
// depend only on the sampled direction): draws (u,v,w), samples the BSDF, fills the ray.
// Returns false when sample_bsdf fails (the reference then returns an EMPTY record).
template <typename R, int MS>
GD_D bool start_iteration(const DScene<R> &sc, const Vertex<R> &vtx, const V3<R> &dir_view, PathRec<R> &pr, RayF &ray) {
    R u0 = pcg_real<R>(pr.rng), u1 = pcg_real<R>(pr.rng), w = pcg_real<R>(pr.rng);
    BsdfSample<R> s;
    if (!bsdf_sample_ms<R, MS>(sc, vtx, dir_view, u0, u1, w, s)) return false;
    if (s.eta != R(0)) pr.eta_scale /= (s.eta * s.eta);
    pr.org = vtx.position; pr.dir = s.dir_out;
    pr.f = bsdf_eval_ms<R, MS>(sc, vtx, dir_view, s.dir_out);
    pr.p2 = bsdf_pdf_ms<R, MS>(sc, vtx, dir_view, s.dir_out);
    ray.ox = (float)pr.org.x; ray.oy = (float)pr.org.y; ray.oz = (float)pr.org.z; ray.tnear = (float)sc.eps;
    ray.dx = (float)pr.dir.x; ray.dy = (float)pr.dir.y; ray.dz = (float)pr.dir.z; ray.tfar = __int_as_float(0x7f800000);
    return true;
}

// ---------------------------------------------------------------- primary stage
template <typename R, int MS, int MINB = 1>
__global__ void __launch_bounds__(128, MINB) k_primary(DScene<R> sc, WaveParams wp, const GenRec<R> *gen, const HitRec *hits,
                                                 PathRec<R> *out, OffRec<R> *off_out, RayF *rays_out, int *out_count) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool alive = false;
    PathRec<R> pr; OffRec<R> off; RayF ray;
    if (i < wp.n_paths) {
        GenRec<R> g; load_soa(g, gen, i, wp.n_paths);
        Hit h0 = load_hit(hits + i);
        if (h0.shape_id >= 0) {   // inactive padding slots carry a ray that cannot hit
            const int ox[5] = {0, -1, 0, 1, 0}, oy[5] = {0, 0, 1, 0, -1};
            const R spread0 = R(0.25) / R(sc.width > sc.height ? sc.width : sc.height);   // init_ray_differential
            V3<R> o, d;
            sample_primary<R>(sc, (R(g.px) + g.rng_x) / R(sc.width), (R(g.py) + g.rng_y) / R(sc.height), o, d);
            Vertex<R> vtx;
            make_vertex_ms<R, MS>(sc, o, d, 0.f, __int_as_float(0x7f800000), h0, R(0), spread0, vtx);
            pr.px = g.px; pr.py = g.py; pr.id = g.id; pr.rng = g.rng; pr.nv = 3;
            pr.thr = V3<R>(1, 1, 1); pr.rad = V3<R>(0, 0, 0); pr.eta_scale = R(1);
            pr.contrib[0] = pr.contrib[1] = pr.contrib[2] = 1.0; pr.prob = 1.0;
            V3<R> dir_view = -d;
            if (is_light(sc, vtx.shape_id)) {
                V3<R> e = emission<R>(sc, vtx, dir_view);
                pr.rad = pr.rad + pr.thr * e;
                pr.contrib[0] = e.x; pr.contrib[1] = e.y; pr.contrib[2] = e.z;
            }
            // offsets: enabled iff they hit something with the base's material (path_tracing.h:424-443);
            // each enabled one replays the base's (u,v,w) at its own vertex (path_tracing.h:774-959)
            const bool loop_runs = sc.max_depth == -1 || 3 <= sc.max_depth + 1;
            Pcg r2 = pr.rng;
            const R u0 = pcg_real<R>(r2), u1 = pcg_real<R>(r2), w = pcg_real<R>(r2);
            off.enabled = 0; off.sampled = 0;
            for (int k = 0; k < 4; k++) {
                off.contrib[k] = V3<R>(1, 1, 1); off.p2s[k] = R(1);
                Hit hk = load_hit(hits + (size_t)(k + 1) * wp.n_paths + i);
                if (hk.shape_id < 0) continue;
                if (sc.shapes[hk.shape_id].material_id != vtx.material_id) continue;
                V3<R> ok_, dk;
                sample_primary<R>(sc, (R(g.px + ox[k + 1]) + g.rng_x) / R(sc.width), (R(g.py + oy[k + 1]) + g.rng_y) / R(sc.height), ok_, dk);
                Vertex<R> vo;
                make_vertex_ms<R, MS>(sc, ok_, dk, 0.f, __int_as_float(0x7f800000), hk, R(0), spread0, vo);
                V3<R> dvo = -dk;
                off.enabled |= 1 << k;
                if (is_light(sc, vo.shape_id)) off.contrib[k] = emission<R>(sc, vo, dvo);
                if (!loop_runs) continue;
                BsdfSample<R> s;
                if (!bsdf_sample_ms<R, MS>(sc, vo, dvo, u0, u1, w, s)) continue;
                R p = bsdf_pdf_ms<R, MS>(sc, vo, dvo, s.dir_out);
                if (p <= R(0)) continue;
                off.p2s[k] = p; off.sampled |= 1 << k;
            }
            if (!loop_runs) {
                finish_with_offsets<R>(wp, pr, &off, -1.0);
            } else {
                if (start_iteration<R, MS>(sc, vtx, dir_view, pr, ray)) alive = true;
                // sample_bsdf failure at the base vertex: `return GraidentPTRadiance{}` -> contributes nothing
            }
        }
    }
    int slot = warp_slot(alive, out_count);
    if (alive) { store_soa(out, slot, wp.n_paths, pr); store_soa(off_out, slot, wp.n_paths, off); store_rec(rays_out + slot, ray); sort_emit(wp, slot, ray); }
}

// ---------------------------------------------------------------- bounce stage
// One path-tracing iteration for the path in `pr` whose sampled ray came back with hit `h`
// (path_tracing.h:746-1010).  Returns true and fills `ray` with the next ray when the path goes on;
// otherwise the path has been accumulated.  `offp` is non-null for the iteration with num_vertices == 3.
template <typename R, int MS>
GD_D bool bounce_step(const DScene<R> &sc, const WaveParams &wp, PathRec<R> &pr, const OffRec<R> *offp, const Hit &h, RayF &ray) {
    bool hit = h.shape_id >= 0;
    Vertex<R> nv;
    R G = R(1);
    if (hit) {
        make_vertex_ms<R, MS>(sc, pr.org, pr.dir, (float)sc.eps, __int_as_float(0x7f800000), h, R(0), R(0), nv);   // intersect(scene, bsdf_ray): default ray differential
        G = fabs(dot(pr.dir, nv.gn)) / distance_squared(nv.position, pr.org);
    }
    if (pr.p2 <= R(0)) { finish_with_offsets<R>(wp, pr, offp, -1.0); return false; }     // path_tracing.h:766-769 (offset jacobians still 1)
    R p2 = pr.p2 * G;
    pr.contrib[0] = pr.contrib[0] * (double)pr.f.x * (double)G; pr.contrib[1] = pr.contrib[1] * (double)pr.f.y * (double)G;
    pr.contrib[2] = pr.contrib[2] * (double)pr.f.z * (double)G;
    pr.prob *= (double)p2;
    if (hit && is_light(sc, nv.shape_id)) {
        V3<R> L = emission<R>(sc, nv, -pr.dir);
        V3<R> C2 = (G * pr.f) * L;
        pr.contrib[0] *= (double)L.x; pr.contrib[1] *= (double)L.y; pr.contrib[2] *= (double)L.z;
        C2 = C2 / p2;
        pr.rad = pr.rad + pr.thr * C2;
    }
    if (!hit) { finish_with_offsets<R>(wp, pr, offp, (double)p2); return false; }
    R rr = R(1);
    if (pr.nv - 1 >= sc.rr_depth) {
        rr = rmin(max3((R(1) / pr.eta_scale) * pr.thr), R(0.95));
        if (pcg_real<R>(pr.rng) > rr) { finish_with_offsets<R>(wp, pr, offp, (double)p2); return false; }
    }
    pr.thr = (pr.thr * (G * pr.f)) / (p2 * rr);
    pr.nv++;
    if (!(sc.max_depth == -1 || pr.nv <= sc.max_depth + 1)) { finish_with_offsets<R>(wp, pr, offp, (double)p2); return false; }
    // the loop goes on: from here every offset is dropped (garbage-vertex rule, path_tracing.h:607-612,1007-1010)
    V3<R> dir_view = -pr.dir;
    return start_iteration<R, MS>(sc, nv, dir_view, pr, ray);
}

// One thread per path of [start, n).  LOOP = false: the launch covers the range (thread i handles path start + i); this is
// the kernel of every ordinary launch — a loop around the body costs it 38 registers (114 -> 152) and a block per SM.
// LOOP = true: grid-stride from `start`; launched after a launch that was sized by an ESTIMATE of n (waves scheduled without
// host read-backs) to take whatever lies beyond the estimate — normally nothing.
template <typename R, int MS, int MINB = 1, bool LOOP = false>
__global__ void __launch_bounds__(128, MINB) k_bounce(DScene<R> sc, WaveParams wp, const PathRec<R> *in, const OffRec<R> *off_in,
                                                const HitRec *hits, const int *in_count, int in_max,
                                                PathRec<R> *out, RayF *rays_out, int *out_count, int start = 0, const int *order = nullptr) {
    const int n = min(*in_count, in_max);
    for (int base = start + blockIdx.x * blockDim.x; base < n; base += gridDim.x * blockDim.x) {
        const int j = base + threadIdx.x;
        bool alive = false;
        PathRec<R> pr; RayF ray;
        if (j < n) {
            // order: the permutation the rays were traced in (origin cell + direction bin): a warp then shades neighbouring
            // hits — the same meshes, textures and, in mixed scenes, mostly the same material
            const int i = order ? __ldg(order + j) : j;
            load_soa(pr, in, i, wp.n_paths);
            const bool first = off_in != nullptr;          // iteration with num_vertices == 3
            OffRec<R> off; if (first) load_soa(off, off_in, i, wp.n_paths);
            Hit h = load_hit(hits + i);
            alive = bounce_step<R, MS>(sc, wp, pr, first ? &off : nullptr, h, ray);
        }
        int slot = warp_slot(alive, out_count);
        if (alive) { store_soa(out, slot, wp.n_paths, pr); store_rec(rays_out + slot, ray); sort_emit(wp, slot, ray); }
        if (!LOOP) break;
    }
}

// ---------------------------------------------------------------- tail
// When only a few paths of a wave are still alive, the per-bounce launches (trace, shade, count
// read-back) cost more than the work in them: this kernel walks each remaining path to its end
// in one launch — trace and shade in the same thread, no queues.
template <typename R, int MS, bool EXACT>
__global__ void __launch_bounds__(TRACE_BLOCK) k_tail(DScene<R> sc, WaveParams wp, DBvh bvh, const PathRec<R> *in, const OffRec<R> *off_in,
                                                      const RayF *rays, int n, const int *n_ptr = nullptr) {
    __shared__ int2 stack[STACK_SMEM * TRACE_BLOCK];
    if (n_ptr) n = min(n, *n_ptr);                      // live count on the device; n is then only an upper bound
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    PathRec<R> pr; RayF ray;
    load_soa(pr, in, i, wp.n_paths); load_rec(ray, rays + i);
    bool first = off_in != nullptr;
    OffRec<R> off; if (first) load_soa(off, off_in, i, wp.n_paths);
    while (true) {
        Hit h; h.t = ray.tfar; h.u = h.v = 0.f; h.shape_id = -1; h.prim_id = -1;
        if (!(ray.tfar < ray.tnear)) {
            const float o[3] = {ray.ox, ray.oy, ray.oz}, d[3] = {ray.dx, ray.dy, ray.dz};
            h = trace_closest<EXACT, false>(bvh, o, d, ray.tnear, ray.tfar, stack);
        }
        bool alive = bounce_step<R, MS>(sc, wp, pr, first ? &off : nullptr, h, ray);
        first = false;
        if (!alive) break;
    }
    }
}

} // namespace gdpt
