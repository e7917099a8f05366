// Wavefront form of the plain `path` integrator — path_render (/root/reference/src/render.cpp:74-117) and
// path_tracing (/root/reference/src/path_tracing.h:13-348): unidirectional path tracing with next-event
// estimation and the power heuristic.  It is the equal-time comparator of the metric and runs on the same
// traversal kernel, BSDFs, textures and queues as gradpath.
//
// One loop iteration of path_tracing at vertex v_i is split at its two ray casts:
//   path_begin_iteration   draws (light_uv, light_w, shape_w), samples the light, builds the shadow ray and the
//                          weighted contribution it would add (:117-171, everything except `occluded`), draws
//                          (bsdf_uv, bsdf_w), samples the BSDF, builds the BSDF ray, evaluates f and p2 (:175-215)
//   [k_trace: shadow ray any-hit, BSDF ray closest hit]
//   path_finish_iteration  adds the light sample if unoccluded (:135-171), G of the BSDF vertex, emission of a
//                          hit area light (:217-232; as shipped the BSDF-sampled light hit is NOT weighted by w2),
//                          Russian roulette, throughput update (:252-265)
// As shipped: a failed sample_bsdf or p2 <= 0 ends the path but keeps its radiance; bounce vertices get the
// default ray differential.  With an environment map (envmap.inl): a primary or BSDF ray that leaves the scene reads
// it (the BSDF-sampled read IS weighted by w2, :233-246), light samples on it cast their shadow ray to infinity, and the
// spread of the ray differential is carried along the path for its footprint (ray.h:50-66).
#pragma once
#include "gdpt_light.cuh"
#include "gdpt_shade.cuh"

namespace gdpt {

template <typename R> struct alignas(16) PathState {
    V3<R> org, dir;          // the BSDF ray in flight (org = current vertex)
    V3<R> f;                 // BSDF*cos towards dir
    R p2;                    // solid-angle pdf of dir (<= 0 also encodes a failed sample_bsdf)
    V3<R> thr, rad;          // current_path_throughput, radiance
    V3<R> nee;               // thr * C1 * w1 of this iteration's light sample; added iff the shadow ray is unoccluded
    R eta_scale;
    R spread;                // ray_diff.spread (its radius stays 0: path_tracing only ever assigns the spread, :196-200)
    R pad0, pad1, pad2;
    Pcg rng;
    int px, py, id, nv;
};

template <typename R>
GD_D void path_finish(const WaveParams &wp, const PathState<R> &ps) {
    if (wp.rec_out) { double *o = wp.rec_out + (size_t)ps.id * 3; o[0] = ps.rad.x; o[1] = ps.rad.y; o[2] = ps.rad.z; return; }
    size_t p = ((size_t)ps.py * wp.width + ps.px) * 3;
    double inv_spp = 1.0 / (double)wp.spp_total;
    double rd[3] = {(double)ps.rad.x, (double)ps.rad.y, (double)ps.rad.z};
    // float mode: a sample whose products left the float range (near-specular chains) is dropped instead of poisoning the
    // pixel; the double mode keeps whatever the reference's arithmetic gives
    if (sizeof(R) == 4 && !(isfinite(rd[0]) && isfinite(rd[1]) && isfinite(rd[2]))) return;
    if (wp.res) {                                    // one 24-byte record per path (zeroed per wave); summed in sample order by k_path_reduce
        double *o = wp.res + (size_t)ps.id * 3;
        for (int c = 0; c < 3; c++) o[c] = rd[c] * inv_spp;
        return;
    }
    for (int c = 0; c < 3; c++) if (rd[c] != 0.0) atomicAdd(wp.img + p + c, rd[c] * inv_spp);
}

// path_render's accumulation (render.cpp:98-113: the samples of a pixel added one after the other) without atomics: thread
// per pixel slot of the wave, the wave's samples added in sample order onto the running image value — bit-reproducible.
static __global__ void __launch_bounds__(256) k_path_reduce(WaveParams wp, const double *res, double *img) {
    const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (p >= wp.npix_pad) return;
    const int tile = (int)(p >> 5), lane = (int)(p & 31);
    const int x = (tile % wp.tiles_x) * 8 + (lane & 7), y = wave_row(wp, tile / wp.tiles_x, lane);
    if (x >= wp.width || y >= wp.y1) return;
    const size_t pix = ((size_t)y * wp.width + x) * 3;
    double a0 = img[pix], a1 = img[pix + 1], a2 = img[pix + 2];
    for (int s = 0; s < wp.n_samples; s++) {
        const double *r = res + ((size_t)s * wp.npix_pad + p) * 3;
        a0 += r[0]; a1 += r[1]; a2 += r[2];
    }
    img[pix] = a0; img[pix + 1] = a1; img[pix + 2] = a2;
}

// power heuristic p_a^2 / (p_a^2 + p_b^2) (p_a > 0).  double: the reference's expression, literally; float: the same
// number from the ratio, because the squares of specular pdfs leave the float range (inf / inf)
template <typename R> GD_D R mis_weight(R pa, R pb) {
    if (sizeof(R) == 8) return (pa * pa) / (pa * pa + pb * pb);
    R r = pb / pa;
    return R(1) / (R(1) + r * r);
}

GD_D void disable_ray(RayF &r) { r.ox = r.oy = r.oz = 0.f; r.dx = r.dy = r.dz = 0.f; r.tnear = 0.f; r.tfar = -1.f; }

// Start of a loop iteration at `vtx`, reached along `dir_in` (the direction of the ray that found it).
// Returns false when the path is over (loop bound reached, or nothing left to trace).
template <typename R, int MS>
GD_D bool path_begin_iteration(const DScene<R> &sc, const Vertex<R> &vtx, const V3<R> &dir_in, PathState<R> &ps, RayF &ray, RayF &shadow) {
    if (!(sc.max_depth == -1 || ps.nv <= sc.max_depth + 1)) return false;
    const V3<R> dir_view = -dir_in;
    // ---- next-event estimation
    R lu0 = pcg_real<R>(ps.rng), lu1 = pcg_real<R>(ps.rng), light_w = pcg_real<R>(ps.rng), shape_w = pcg_real<R>(ps.rng);
    ps.nee = V3<R>(R(0), R(0), R(0));
    disable_ray(shadow);
    if (sc.n_lights > 0) {
        int light_id = sample_light<R>(sc, light_w);
        LightPoint<R> pl = sample_point_on_light<R>(sc, light_id, vtx.position, lu0, lu1, shape_w);
        const bool env = is_envmap(sc, light_id);
        V3<R> dir_light = env ? -pl.normal : normalize(pl.position - vtx.position);
        // G of the light sample as if unoccluded (:128-149); the occlusion test itself is the shadow ray
        R G = env ? R(1) : rmax(-dot(dir_light, pl.normal), R(0)) / distance_squared(pl.position, vtx.position);
        R p1 = light_pmf<R>(sc, light_id) * pdf_point_on_light<R>(sc, light_id, pl, vtx.position);
        if (G > R(0) && p1 > R(0)) {
            V3<R> fl = bsdf_eval_ms<R, MS>(sc, vtx, dir_view, dir_light);
            V3<R> L = light_emission<R>(sc, light_id, -dir_light, pl);
            V3<R> C1 = (G * fl) * L;
            R p2l = bsdf_pdf_ms<R, MS>(sc, vtx, dir_view, dir_light) * G;
            R w1 = mis_weight<R>(p1, p2l);
            C1 = C1 / p1;
            ps.nee = (ps.thr * C1) * w1;
            if (ps.nee.x != R(0) || ps.nee.y != R(0) || ps.nee.z != R(0)) {
                shadow.ox = (float)vtx.position.x; shadow.oy = (float)vtx.position.y; shadow.oz = (float)vtx.position.z;
                shadow.dx = (float)dir_light.x; shadow.dy = (float)dir_light.y; shadow.dz = (float)dir_light.z;
                shadow.tnear = (float)sc.eps;
                shadow.tfar = env ? __int_as_float(0x7f800000) : (float)((R(1) - sc.eps) * distance(pl.position, vtx.position));
            }
        }
    }
    // ---- BSDF sampling
    R u0 = pcg_real<R>(ps.rng), u1 = pcg_real<R>(ps.rng), w = pcg_real<R>(ps.rng);
    BsdfSample<R> s;
    disable_ray(ray);
    ps.org = vtx.position; ps.p2 = R(-1); ps.f = V3<R>(R(0), R(0), R(0)); ps.dir = V3<R>(R(0), R(0), R(1));
    if (bsdf_sample_ms<R, MS>(sc, vtx, dir_view, u0, u1, w, s)) {
        if (s.eta == R(0)) ps.spread = rd_reflect<R>(R(0), ps.spread, vtx.mean_curvature, s.roughness);
        else { ps.spread = rd_refract<R>(R(0), ps.spread, vtx.mean_curvature, s.eta, s.roughness); ps.eta_scale /= (s.eta * s.eta); }
        ps.dir = s.dir_out;
        ps.f = bsdf_eval_ms<R, MS>(sc, vtx, dir_view, s.dir_out);
        ps.p2 = bsdf_pdf_ms<R, MS>(sc, vtx, dir_view, s.dir_out);
        if (ps.p2 > R(0)) {   // with p2 <= 0 the reference still casts the ray, but nothing depends on its result
            ray.ox = (float)ps.org.x; ray.oy = (float)ps.org.y; ray.oz = (float)ps.org.z; ray.tnear = (float)sc.eps;
            ray.dx = (float)ps.dir.x; ray.dy = (float)ps.dir.y; ray.dz = (float)ps.dir.z; ray.tfar = __int_as_float(0x7f800000);
        }
    }
    // nothing to wait for: the light sample adds nothing and the path cannot go on
    if (shadow.tfar < shadow.tnear && !(ps.p2 > R(0))) return false;
    return true;
}

// Second half of the iteration, once both rays are back.  Returns false when the path is over.
template <typename R, int MS>
GD_D bool path_finish_iteration(const DScene<R> &sc, PathState<R> &ps, bool shadow_traced, bool occluded, const Hit &h, Vertex<R> &nv) {
    if (shadow_traced && !occluded) ps.rad = ps.rad + ps.nee;
    if (!(ps.p2 > R(0))) return false;                                          // sample_bsdf failed (:190-192) or p2 <= 0 (:213-215)
    const bool hit = h.shape_id >= 0;
    R G = R(1);
    if (hit) {
        make_vertex_ms<R, MS>(sc, ps.org, ps.dir, (float)sc.eps, __int_as_float(0x7f800000), h, R(0), R(0), nv);
        G = fabs(dot(ps.dir, nv.gn)) / distance_squared(nv.position, ps.org);
    }
    R p2 = ps.p2 * G;
    if (hit && is_light(sc, nv.shape_id)) {
        V3<R> L = emission<R>(sc, nv, -ps.dir);
        V3<R> C2 = (G * ps.f) * L;
        C2 = C2 / p2;
        ps.rad = ps.rad + ps.thr * C2;
    }
    if (!hit && sc.env_light >= 0) {             // :233-246: the environment map seen along the BSDF ray, weighted by w2
        V3<R> L = env_emission<R>(sc, -ps.dir, ps.spread);
        V3<R> C2 = (G * ps.f) * L;
        LightPoint<R> lp; lp.position = V3<R>(R(0), R(0), R(0)); lp.normal = -ps.dir;
        R p1 = light_pmf<R>(sc, sc.env_light) * env_pdf_point<R>(sc, lp);
        R w2 = mis_weight<R>(p2, p1);
        C2 = C2 / p2;
        ps.rad = ps.rad + (ps.thr * C2) * w2;
    }
    if (!hit) return false;
    R rr = R(1);
    if (ps.nv - 1 >= sc.rr_depth) {
        rr = rmin(max3((R(1) / ps.eta_scale) * ps.thr), R(0.95));
        if (pcg_real<R>(ps.rng) > rr) return false;
    }
    ps.thr = (ps.thr * (G * ps.f)) / (p2 * rr);
    ps.nv++;
    return true;
}

// ---------------------------------------------------------------- kernels
template <typename R>
__global__ void __launch_bounds__(128) k_path_generate(DScene<R> sc, WaveParams wp, PathState<R> *out, RayF *rays) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= wp.n_paths) return;
    PathState<R> ps; memset(&ps, 0, sizeof(ps));
    ps.id = i; ps.px = -1; ps.py = -1;
    unsigned long long stream = 0; bool active = false;
    if (wp.xs) { ps.px = wp.xs[i]; ps.py = wp.ys[i]; stream = wp.streams[i]; active = true; }
    else {
        long long s = i / wp.npix_pad, p = i % wp.npix_pad;
        int tile = (int)(p >> 5), lane = (int)(p & 31);
        int x = (tile % wp.tiles_x) * 8 + (lane & 7), y = wave_row(wp, tile / wp.tiles_x, lane);
        if (x < wp.width && y < wp.y1) {
            ps.px = x; ps.py = y; active = true;
            stream = (unsigned long long)(wp.sample0 + s) * (unsigned long long)wp.width * wp.height + (unsigned long long)y * wp.width + x;
        }
    }
    RayF r; disable_ray(r);
    if (active) {
        ps.rng = pcg_init(stream, wp.seed);
        // path_tracing.h:21-22 draws both numbers inside ONE constructor call, `Vector2 screen_pos((x + next(rng)) / w,
        // (y + next(rng)) / h)`: the order is unspecified in C++, and gcc (the oracle build) evaluates the arguments
        // right to left — the first draw goes to y.  (grad_path_tracing draws rng_x, rng_y in separate statements.)
        R ry = pcg_real<R>(ps.rng), rx = pcg_real<R>(ps.rng);
        sample_primary<R>(sc, (R(ps.px) + rx) / R(sc.width), (R(ps.py) + ry) / R(sc.height), ps.org, ps.dir);
        ps.thr = V3<R>(R(1), R(1), R(1)); ps.eta_scale = R(1); ps.nv = 3;
        ps.spread = R(0.25) / R(sc.width > sc.height ? sc.width : sc.height);   // init_ray_differential (ray.h:35-37)
        r.ox = (float)ps.org.x; r.oy = (float)ps.org.y; r.oz = (float)ps.org.z; r.dx = (float)ps.dir.x; r.dy = (float)ps.dir.y; r.dz = (float)ps.dir.z;
        r.tnear = 0.f; r.tfar = __int_as_float(0x7f800000);
    }
    store_soa(out, i, wp.n_paths, ps);
    store_rec(rays + i, r);
}

// FIRST: the hit is the primary hit (path_tracing.h:23-81); otherwise the BSDF-ray hit of the iteration in flight.
template <typename R, int MS, bool FIRST>
__global__ void __launch_bounds__(128) k_path_shade(DScene<R> sc, WaveParams wp, const PathState<R> *in, const HitRec *hits, const RayF *shadow_in,
                                                    const HitRec *shadow_hits, const int *in_count, int in_max,
                                                    PathState<R> *out, RayF *rays_out, RayF *shadow_out, int *out_count, const int *order = nullptr) {
    int n = FIRST ? in_max : min(*in_count, in_max);
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    // order: the permutation the rays were traced in — generic-material scenes shade in that order so that a warp meets one
    // material (gdpt_ctx::shade_sorted, as in k_bounce)
    const int i = (order && j < n) ? __ldg(order + j) : j;
    bool alive = false;
    PathState<R> ps; RayF ray, shadow;
    if (j < n) {
        load_soa(ps, in, i, wp.n_paths);
        Hit h = load_hit(hits + i);
        if (FIRST) {
            if (ps.px >= 0) {
                if (h.shape_id >= 0) {
                    Vertex<R> vtx;
                    make_vertex_ms<R, MS>(sc, ps.org, ps.dir, 0.f, __int_as_float(0x7f800000), h, R(0), ps.spread, vtx);
                    if (is_light(sc, vtx.shape_id)) ps.rad = ps.rad + ps.thr * emission<R>(sc, vtx, -ps.dir);
                    alive = path_begin_iteration<R, MS>(sc, vtx, ps.dir, ps, ray, shadow);
                } else if (sc.env_light >= 0) {
                    ps.rad = env_emission<R>(sc, -ps.dir, ps.spread);                             // :28-38: the background
                }
                if (!alive) path_finish<R>(wp, ps);
            }
        } else {
            const float4 sa = __ldg(reinterpret_cast<const float4 *>(shadow_in + i));       // origin, tnear
            const float4 sb = __ldg(reinterpret_cast<const float4 *>(shadow_in + i) + 1);   // direction, tfar
            const bool shadow_traced = !(sb.w < sa.w);
            const bool occluded = shadow_traced && load_hit(shadow_hits + i).shape_id >= 0;
            Vertex<R> nv;
            V3<R> dir_in = ps.dir;
            if (path_finish_iteration<R, MS>(sc, ps, shadow_traced, occluded, h, nv)) alive = path_begin_iteration<R, MS>(sc, nv, dir_in, ps, ray, shadow);
            if (!alive) path_finish<R>(wp, ps);
        }
    }
    int slot = warp_slot(alive, out_count);
    if (alive) { store_soa(out, slot, wp.n_paths, ps); store_rec(rays_out + slot, ray); store_rec(shadow_out + slot, shadow); sort_emit(wp, slot, ray); }
}

// Last paths of a wave: walked to their end in one launch (both ray casts and the shading in the same thread), like
// k_tail does for gradpath — the per-bounce launches and count read-backs cost more than the work left in them.
template <typename R, int MS, bool EXACT>
__global__ void __launch_bounds__(TRACE_BLOCK) k_path_tail(DScene<R> sc, WaveParams wp, DBvh bvh, const PathState<R> *in, const RayF *rays, const RayF *shadows, int n) {
    __shared__ int2 stack[STACK_SMEM * TRACE_BLOCK];
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    PathState<R> ps; RayF ray, shadow;
    load_soa(ps, in, i, wp.n_paths); load_rec(ray, rays + i); load_rec(shadow, shadows + i);
    while (true) {
        Hit h; h.t = ray.tfar; h.u = h.v = 0.f; h.shape_id = -1; h.prim_id = -1;
        if (!(ray.tfar < ray.tnear)) {
            const float o[3] = {ray.ox, ray.oy, ray.oz}, d[3] = {ray.dx, ray.dy, ray.dz};
            h = trace_closest<EXACT, false>(bvh, o, d, ray.tnear, ray.tfar, stack);
        }
        const bool shadow_traced = !(shadow.tfar < shadow.tnear);
        bool occluded = false;
        if (shadow_traced) {
            const float o[3] = {shadow.ox, shadow.oy, shadow.oz}, d[3] = {shadow.dx, shadow.dy, shadow.dz};
            occluded = trace_occluded<EXACT>(bvh, o, d, shadow.tnear, shadow.tfar, stack);
        }
        Vertex<R> nv;
        V3<R> dir_in = ps.dir;
        bool alive = path_finish_iteration<R, MS>(sc, ps, shadow_traced, occluded, h, nv);
        if (alive) alive = path_begin_iteration<R, MS>(sc, nv, dir_in, ps, ray, shadow);
        if (!alive) break;
    }
    path_finish<R>(wp, ps);
}

} // namespace gdpt
