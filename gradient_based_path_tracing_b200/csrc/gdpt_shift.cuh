// Shift-mapped gradient sampling (SURVEY §7.3 "smallgdpt" mode, §8(f) N2): the algorithm the reference's
// grad_path_tracing was meant to be (small_gdpt.cpp:153-352, restated for lajolla scenes), opt-in through
// gdpt_render_params.offset_mode = GDPT_OFFSETS_SHIFT.  The default mode stays the as-shipped semantics.
//
//   base path    BSDF sampling only; it counts only when it ENDS ON AN AREA LIGHT (small_gdpt.cpp:153-197).  C = prod f_j * Le,
//                P = prod p_j, both in the solid-angle measure (lajolla's eval carries the cosine).
//   offset path  same pixel jitter one pixel over, and per vertex the base's (u, v, w) replayed — J *= p_base / p_offset
//                (:274-280) — until the first base vertex b_d whose successor b_{d+1} is also diffuse (or the light):
//                there the offset reconnects to b_{d+1} (visibility ray, :248-270),
//                J *= (cos'/d'^2) / (cos/d^2) with the cosines at b_{d+1}, re-evaluates the BSDF of b_{d+1} for its new
//                incoming direction and shares the rest of the base path.  Every offset vertex must be of the base
//                vertex's class (diffuse / glossy / light, :232-237), and the offset must end on the light as well.
//   estimators   r += C/P, d += +-(C - C_s J) w / P with w = P / (P + P_s J) (w = 1, C_s = 0 for a failed shift): the
//                reference's accumulation lines (render.cpp:311-327) fed with consistent inputs.  X0 = (x-1, y),
//                X1 = (x+1, y); the y pair is (x, y-1) / (x, y+1) here, so that cy0(x,y) + cy1(x,y-1) are two estimates
//                of the same backward difference (the reference's pairing assumes small_gdpt's flipped rows, which
//                lajolla does not have: SURVEY §3.2).
// Russian roulette starts at the scene's rr_depth with a CONSTANT survival probability: a throughput-dependent one would
// differ between a path and its shift (the offset path, taken as a base path, would have survived with another
// probability), the two MIS weights of a pixel pair would no longer sum to one and the gradients would be biased.  A
// constant multiplies base and offset alike: it cancels in w and scales both contributions.
//
// One thread walks one pixel-sample: base path and four offsets, ray casts inline — the offsets need the base path's
// vertices and random numbers, which would otherwise travel through HBM per bounce.  Two forms of the same arithmetic:
// k_shift (program order, kept as the readable statement and the A/B baseline) and k_shift_sm (per-lane state machine
// with one ray-cast site and persistent lanes, the default).
#pragma once
#include "gdpt_wave.cuh"

namespace gdpt {

constexpr int SHIFT_MAX_VERTS = 24;     // surface vertices of a base path (small_gdpt.cpp MAX_DEPTH 32)
constexpr double SHIFT_RR_SURVIVAL = 0.75;
constexpr int SHIFT_MAX_PREFIX = 12;    // replayed vertices before the reconnection; longer glossy chains keep their primal, lose their offsets

template <typename R> GD_D bool shift_is_diffuse(const DScene<R> &sc, int material_id) {
    const int t = sc.materials[material_id].type;
    return t == MAT_LAMBERTIAN || t == MAT_DISNEY_DIFFUSE;
}
template <typename R> GD_D int shift_class(const DScene<R> &sc, const Vertex<R> &v) {
    const bool light = is_light(sc, v.shape_id);
    return (light || shift_is_diffuse(sc, v.material_id) ? 1 : 0) | (light ? 2 : 0);
}

template <typename R, bool EXACT>
GD_D Hit shift_cast(const DBvh &bvh, const V3<R> &o, const V3<R> &d, float tnear, int2 *stack, unsigned &n_rays) {
    const float of[3] = {(float)o.x, (float)o.y, (float)o.z}, df[3] = {(float)d.x, (float)d.y, (float)d.z};
    n_rays++;
    return trace_closest<EXACT, false>(bvh, of, df, tnear, __int_as_float(0x7f800000), stack);
}

// Accumulation of one sample (render.cpp:311-327) with the image border the screened-Poisson solve assumes: fourierSolve
// reads gx(0,y) as u(0,y) - u(1,y) and gy(x,0) as u(x,0) - u(x,1) (whole-sample symmetric extension, render.cpp:216-224), so
// column 0 / row 0 of the gradient images receive MINUS the estimates of column 1 / row 1 instead of half an estimate
// across the film's edge (the as-shipped mode keeps the latter).  k_merge then leaves them as they are (x == 0: cx = cx0).
template <typename R>
GD_D void finish_shift(const WaveParams &wp, int px, int py, int id, const V3<R> &rad, const double contrib[3], double prob,
                       const double cS[4][3], const double wS[4]) {
    if (wp.rec_out) { finish_path<R>(wp, px, py, id, rad, contrib, prob, cS, wS); return; }
    if (!(prob > 0.0)) return;
    const size_t p = ((size_t)py * wp.width + px) * 3, row = (size_t)wp.width * 3;
    const double spp = (double)wp.spp_total, inv_spp = 1.0 / spp;
    const double rd[3] = {(double)rad.x, (double)rad.y, (double)rad.z};
    if (wp.res) {            // one 120-byte record per path (zeroed per wave): r, vx0, vy0, vx1, vy1; the border rule is k_shift_reduce's
        double *o = wp.res + (size_t)id * 15;
        for (int c = 0; c < 3; c++) {
            o[c] = rd[c] * inv_spp;
            o[3 + c] = (contrib[c] - cS[0][c]) * (wS[0] / (prob * spp)); o[6 + c] = (contrib[c] - cS[1][c]) * (wS[1] / (prob * spp));
            o[9 + c] = (cS[2][c] - contrib[c]) * (wS[2] / (prob * spp)); o[12 + c] = (cS[3][c] - contrib[c]) * (wS[3] / (prob * spp));
        }
        return;
    }
    for (int c = 0; c < 3; c++) {
        if (rd[c] != 0.0) atomicAdd(wp.img + p + c, rd[c] * inv_spp);
        const double vx0 = (contrib[c] - cS[0][c]) * (wS[0] / (prob * spp)), vy0 = (contrib[c] - cS[1][c]) * (wS[1] / (prob * spp));
        const double vx1 = (cS[2][c] - contrib[c]) * (wS[2] / (prob * spp)), vy1 = (cS[3][c] - contrib[c]) * (wS[3] / (prob * spp));
        if (px > 0 && vx0 != 0.0) { atomicAdd(wp.cx0 + p + c, vx0); if (px == 1) atomicAdd(wp.cx0 + p - 3 + c, -vx0); }
        if (py > 0 && vy0 != 0.0) { atomicAdd(wp.cy0 + p + c, vy0); if (py == 1) atomicAdd(wp.cy0 + p - row + c, -vy0); }
        if (px < wp.width - 1 && vx1 != 0.0) { atomicAdd(wp.cx1 + p + c, vx1); if (px == 0) atomicAdd(wp.cx0 + p + c, -vx1); }
        if (py < wp.height - 1 && vy1 != 0.0) { atomicAdd(wp.cy1 + p + c, vy1); if (py == 0) atomicAdd(wp.cy0 + p + c, -vy1); }
    }
}

// finish_shift's accumulation without atomics: thread per pixel slot of the wave, the wave's samples in sample order.  A pixel of
// column 0 / row 0 also takes, sample by sample after its own terms, minus the X0 / Y0 estimate of its neighbour in column 1 /
// row 1 (the border rule above): that neighbour lies in the same 8x4 tile, one lane / one tile row further.
static __global__ void __launch_bounds__(256) k_shift_reduce(WaveParams wp, const double *res, double *img, double *cx0, double *cy0, double *cx1, double *cy1) {
    const long long p = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (p >= wp.npix_pad) return;
    const int tile = (int)(p >> 5), lane = (int)(p & 31);
    const int x = (tile % wp.tiles_x) * 8 + (lane & 7), y = wave_row(wp, tile / wp.tiles_x, lane);
    if (x >= wp.width || y >= wp.y1) return;
    const size_t pix = ((size_t)y * wp.width + x) * 3;
    const bool has_l = x > 0, has_u = y > 0, has_r = x < wp.width - 1, has_d = y < wp.height - 1;
    const bool nb_x = x == 0 && wp.width > 1, nb_y = y == 0 && wp.height > 1 && wp.y1 > 1;      // the neighbour's record exists in this wave
    double a[15];
    double *bufs[5] = {img, cx0, cy0, cx1, cy1};
#pragma unroll
    for (int k = 0; k < 15; k++) a[k] = bufs[k / 3][pix + k % 3];
    for (int s = 0; s < wp.n_samples; s++) {
        const double *r = res + ((size_t)s * wp.npix_pad + p) * 15;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            a[c] += r[c];
            if (has_l) a[3 + c] += r[3 + c];
            if (has_u) a[6 + c] += r[6 + c];
            if (has_r) { a[9 + c] += r[9 + c]; if (x == 0) a[3 + c] += -r[9 + c]; }
            if (has_d) { a[12 + c] += r[12 + c]; if (y == 0) a[6 + c] += -r[12 + c]; }
            if (nb_x) a[3 + c] += -r[15 + 3 + c];               // pixel (1, y): next lane of the tile
            if (nb_y) a[6 + c] += -r[15 * 8 + 6 + c];           // pixel (x, 1): next row of the tile
        }
    }
    if (y == 1 && wp.y0 == 1) {          // a row band that starts at row 1: row 0 belongs to another call, so this thread carries the border term there
        double b[3] = {cy0[pix - (size_t)wp.width * 3], cy0[pix - (size_t)wp.width * 3 + 1], cy0[pix - (size_t)wp.width * 3 + 2]};
        for (int s = 0; s < wp.n_samples; s++) {
            const double *r = res + ((size_t)s * wp.npix_pad + p) * 15;
            for (int c = 0; c < 3; c++) b[c] += -r[6 + c];
        }
        for (int c = 0; c < 3; c++) cy0[pix - (size_t)wp.width * 3 + c] = b[c];
    }
#pragma unroll
    for (int k = 0; k < 15; k++) bufs[k / 3][pix + k % 3] = a[k];
}

template <typename R, int MS, bool EXACT>
__global__ void __launch_bounds__(TRACE_BLOCK) k_shift(DScene<R> sc, WaveParams wp, DBvh bvh, unsigned long long *ray_counter) {
    __shared__ int2 stack[STACK_SMEM * TRACE_BLOCK];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned n_rays = 0;
    do {
        if (i >= wp.n_paths) break;
        int px, py; unsigned long long stream;
        if (wp.xs) { px = wp.xs[i]; py = wp.ys[i]; stream = wp.streams[i]; }
        else {
            long long s = i / wp.npix_pad, p = i % wp.npix_pad;
            int tile = (int)(p >> 5), lane = (int)(p & 31);
            px = (tile % wp.tiles_x) * 8 + (lane & 7); py = wave_row(wp, tile / wp.tiles_x, lane);
            if (px >= wp.width || py >= wp.y1) break;
            stream = (unsigned long long)(wp.sample0 + s) * (unsigned long long)wp.width * wp.height + (unsigned long long)py * wp.width + px;
        }
        double cS[4][3], wS[4], contrib[3] = {0, 0, 0};
        for (int k = 0; k < 4; k++) { cS[k][0] = cS[k][1] = cS[k][2] = 0.0; wS[k] = 1.0; }
        V3<R> rad(0, 0, 0);
        const float inf = __int_as_float(0x7f800000);
        const R spread0 = R(0.25) / R(sc.width > sc.height ? sc.width : sc.height);
        Pcg rng = pcg_init(stream, wp.seed);
        const R rx = pcg_real<R>(rng), ry = pcg_real<R>(rng);

        // ------------------------------------------------------------ base path
        R pu0[SHIFT_MAX_PREFIX], pu1[SHIFT_MAX_PREFIX], pw[SHIFT_MAX_PREFIX], pp[SHIFT_MAX_PREFIX];
        unsigned char pclass[SHIFT_MAX_PREFIX + 1];
        V3<R> o, d;
        sample_primary<R>(sc, (R(px) + rx) / R(sc.width), (R(py) + ry) / R(sc.height), o, d);
        Hit h = shift_cast<R, EXACT>(bvh, o, d, 0.f, stack, n_rays);
        bool base_ok = false, offsets_ok = true;
        int L = -1, dr = -1;
        double C[3] = {1, 1, 1}, P = 1, inv_rr = 1, Sf[3] = {1, 1, 1}, Sp = 1;
        V3<R> Le(0, 0, 0), wr(0, 0, 1), pos_dr(0, 0, 0);
        Vertex<R> vr;
        if (h.shape_id >= 0) {
            Vertex<R> v;
            make_vertex_ms<R, MS>(sc, o, d, 0.f, inf, h, R(0), spread0, v);
            V3<R> dir_in = d, prev_pos = o;
            bool prev_diffuse = false;
            for (int j = 0; j < SHIFT_MAX_VERTS; j++) {
                const int cls = shift_class(sc, v);
                if (dr < 0 && j > 0 && prev_diffuse && (cls & 1)) { dr = j - 1; vr = v; pos_dr = prev_pos; }
                if (dr < 0 || j <= dr) { if (j <= SHIFT_MAX_PREFIX) pclass[j] = (unsigned char)cls; else offsets_ok = false; }
                if (cls & 2) { Le = emission<R>(sc, v, -dir_in); L = j; base_ok = true; break; }
                if (sc.max_depth != -1 && j + 2 > sc.max_depth) break;        // path_tracing.h's bound on num_vertices (camera, v0 .. v_{j+1})
                const R u0 = pcg_real<R>(rng), u1 = pcg_real<R>(rng), w = pcg_real<R>(rng);
                BsdfSample<R> s;
                if (!bsdf_sample_ms<R, MS>(sc, v, -dir_in, u0, u1, w, s)) break;
                const V3<R> f = bsdf_eval_ms<R, MS>(sc, v, -dir_in, s.dir_out);
                const R p = bsdf_pdf_ms<R, MS>(sc, v, -dir_in, s.dir_out);
                if (!(p > R(0)) || !(f.x > R(0) || f.y > R(0) || f.z > R(0))) break;
                if (dr < 0) { if (j < SHIFT_MAX_PREFIX) { pu0[j] = u0; pu1[j] = u1; pw[j] = w; pp[j] = p; } else offsets_ok = false; }
                else if (j == dr + 1) wr = s.dir_out;
                else { Sf[0] *= (double)f.x; Sf[1] *= (double)f.y; Sf[2] *= (double)f.z; Sp *= (double)p; }
                C[0] *= (double)f.x; C[1] *= (double)f.y; C[2] *= (double)f.z; P *= (double)p;
                if (j + 2 >= sc.rr_depth) {
                    if (!(pcg_real<R>(rng) <= R(SHIFT_RR_SURVIVAL))) break;
                    inv_rr /= SHIFT_RR_SURVIVAL;
                }
                h = shift_cast<R, EXACT>(bvh, v.position, s.dir_out, (float)sc.eps, stack, n_rays);
                if (h.shape_id < 0) break;
                prev_pos = v.position; prev_diffuse = (cls & 1) != 0; dir_in = s.dir_out;
                const V3<R> org = v.position;
                make_vertex_ms<R, MS>(sc, org, dir_in, (float)sc.eps, inf, h, R(0), R(0), v);
            }
        }
        if (!base_ok) { if (wp.rec_out) finish_path<R>(wp, px, py, i, rad, contrib, 1.0, cS, wS); break; }
        for (int c = 0; c < 3; c++) contrib[c] = C[c] * (double)(c == 0 ? Le.x : c == 1 ? Le.y : Le.z) * inv_rr;
        rad = V3<R>((R)(contrib[0] / P), (R)(contrib[1] / P), (R)(contrib[2] / P));

        // ------------------------------------------------------------ the four offsets
        const int ox[4] = {-1, 0, 1, 0}, oy[4] = {0, -1, 0, 1};      // X0, Y0, X1, Y1 (see the header for the y pair)
        const int jend = dr >= 0 ? dr : L;
        for (int k = 0; k < 4 && offsets_ok; k++) {
            // a neighbour outside the film has no pixel to pair with (image mode; records keep all four)
            if (!wp.rec_out && (px + ox[k] < 0 || px + ox[k] >= wp.width || py + oy[k] < 0 || py + oy[k] >= wp.height)) continue;
            V3<R> ok_, dk;
            sample_primary<R>(sc, (R(px + ox[k]) + rx) / R(sc.width), (R(py + oy[k]) + ry) / R(sc.height), ok_, dk);
            Hit ho = shift_cast<R, EXACT>(bvh, ok_, dk, 0.f, stack, n_rays);
            if (ho.shape_id < 0) continue;
            Vertex<R> vo;
            make_vertex_ms<R, MS>(sc, ok_, dk, 0.f, inf, ho, R(0), spread0, vo);
            V3<R> din = dk;
            double Cs[3] = {1, 1, 1}, Ps = 1, J = 1;
            bool ok = true;
            for (int j = 0; ; j++) {
                if (shift_class(sc, vo) != (int)pclass[j]) { ok = false; break; }
                if (j == jend) break;
                BsdfSample<R> s;
                if (!bsdf_sample_ms<R, MS>(sc, vo, -din, pu0[j], pu1[j], pw[j], s)) { ok = false; break; }
                const V3<R> f = bsdf_eval_ms<R, MS>(sc, vo, -din, s.dir_out);
                const R p = bsdf_pdf_ms<R, MS>(sc, vo, -din, s.dir_out);
                if (!(p > R(0))) { ok = false; break; }
                Cs[0] *= (double)f.x; Cs[1] *= (double)f.y; Cs[2] *= (double)f.z; Ps *= (double)p; J *= (double)pp[j] / (double)p;
                ho = shift_cast<R, EXACT>(bvh, vo.position, s.dir_out, (float)sc.eps, stack, n_rays);
                if (ho.shape_id < 0) { ok = false; break; }
                din = s.dir_out;
                const V3<R> org = vo.position;
                make_vertex_ms<R, MS>(sc, org, din, (float)sc.eps, inf, ho, R(0), R(0), vo);
            }
            if (!ok) continue;
            if (dr < 0) {                                    // replayed all the way: the offset's own light vertex
                const V3<R> e = emission<R>(sc, vo, -din);
                Cs[0] *= (double)e.x; Cs[1] *= (double)e.y; Cs[2] *= (double)e.z;
            } else {                                         // reconnect vo -> b_{dr+1}
                const V3<R> dv = vr.position - vo.position;
                const R d2s = dot(dv, dv);
                if (!(d2s > R(0))) continue;
                const R dist = sqrt(d2s);
                const V3<R> ds = dv / dist;
                const float of[3] = {(float)vo.position.x, (float)vo.position.y, (float)vo.position.z}, df[3] = {(float)ds.x, (float)ds.y, (float)ds.z};
                n_rays++;
                if (trace_occluded<EXACT>(bvh, of, df, (float)sc.eps, (float)((R(1) - sc.eps) * dist), stack)) continue;
                const V3<R> f = bsdf_eval_ms<R, MS>(sc, vo, -din, ds);
                const R p = bsdf_pdf_ms<R, MS>(sc, vo, -din, ds);
                if (!(p > R(0))) continue;
                const V3<R> bv = vr.position - pos_dr;
                const R d2b = dot(bv, bv);
                const R cos_s = fabs(dot(ds, vr.gn)), cos_b = fabs(dot(bv, vr.gn)) / sqrt(d2b);
                if (!(cos_b > R(0)) || !(d2b > R(0))) continue;
                J *= ((double)cos_s / (double)d2s) / ((double)cos_b / (double)d2b);
                Cs[0] *= (double)f.x; Cs[1] *= (double)f.y; Cs[2] *= (double)f.z; Ps *= (double)p;
                if (L == dr + 1) {
                    const V3<R> e = emission<R>(sc, vr, -ds);
                    Cs[0] *= (double)e.x; Cs[1] *= (double)e.y; Cs[2] *= (double)e.z;
                } else {
                    const V3<R> ft = bsdf_eval_ms<R, MS>(sc, vr, -ds, wr);
                    const R pt = bsdf_pdf_ms<R, MS>(sc, vr, -ds, wr);
                    if (!(pt > R(0))) continue;
                    Cs[0] *= (double)ft.x * Sf[0] * (double)Le.x; Cs[1] *= (double)ft.y * Sf[1] * (double)Le.y; Cs[2] *= (double)ft.z * Sf[2] * (double)Le.z;
                    Ps *= (double)pt * Sp;
                }
            }
            if (!(J > 0.0) || !isfinite(J) || !isfinite(Ps)) continue;
            for (int c = 0; c < 3; c++) cS[k][c] = Cs[c] * J * inv_rr;
            wS[k] = P / (P + Ps * J);
        }
        finish_shift<R>(wp, px, py, i, rad, contrib, P, cS, wS);
    } while (false);
    // rays traced, one atomic per warp
    for (int s = 16; s > 0; s >>= 1) n_rays += __shfl_xor_sync(0xffffffffu, n_rays, s);
    if ((threadIdx.x & 31) == 0 && n_rays) atomicAdd(ray_counter, (unsigned long long)n_rays);
}

// ---------------------------------------------------------------- the same algorithm as a per-lane state machine
// k_shift above walks base path and offsets in program order: lanes of a warp leave the base loop at different bounces
// and then sit in four different ray-cast call sites, and a warp lasts as long as its longest path — ncu shows 6.2 of
// 32 lanes active per instruction.  Here every lane carries a phase (fetch / base / offset / reconnect), ALL ray casts go
// through one call site per trip of the loop, and a lane that finishes its pixel-sample fetches the next one, so the
// traversal — most of the time — runs with whatever lanes have a ray, whichever stage they are in.  Same arithmetic in
// the same order: records are bit-identical to k_shift's (tests/test_shift_gpu.py).
enum { SH_FETCH = 0, SH_BASE = 1, SH_OFFSET = 2, SH_RECONNECT = 3 };

template <bool EXACT>
GD_D Hit shift_trace(const DBvh &bvh, const float o[3], const float d[3], float tnear, float tfar, bool any, int2 *smem_stack) {
    RayState r; r.init(make_float4(o[0], o[1], o[2], tnear), make_float4(d[0], d[1], d[2], tfar));
    if (bvh.n_nodes == 0) return r.hit;
    TraceStack st; st.smem = smem_stack;
    int sp = 0, cur = 0;
    while (cur != TRACE_DONE) {
        while (cur >= 0) cur = trace_node<0, 0, false, 0, false, 1>(bvh, cur, r, st, sp);
        if (cur != TRACE_DONE) cur = trace_leaf<EXACT>(bvh, cur, r, st, sp, nullptr);
        if (any && r.hit.shape_id >= 0) break;
    }
    return r.hit;
}

template <typename R, int MS, bool EXACT, int MINB>
__global__ void __launch_bounds__(TRACE_BLOCK, MINB) k_shift_sm(DScene<R> sc, WaveParams wp, DBvh bvh, unsigned long long *ray_counter, int *next_sample) {
    __shared__ int2 stack[STACK_SMEM * TRACE_BLOCK];
    const float inf = __int_as_float(0x7f800000);
    const R spread0 = R(0.25) / R(sc.width > sc.height ? sc.width : sc.height);
    const int ox[4] = {-1, 0, 1, 0}, oy[4] = {0, -1, 0, 1};
    unsigned n_rays = 0;
    int phase = SH_FETCH;
    // the sample
    int i = 0, px = 0, py = 0; R rx = R(0), ry = R(0); Pcg rng;
    // the ray in flight
    V3<R> ro(0, 0, 0), rd(0, 0, 1); float rtnear = 0.f, rtfar = inf; bool rany = false;
    // base path
    R pu0[SHIFT_MAX_PREFIX], pu1[SHIFT_MAX_PREFIX], pw[SHIFT_MAX_PREFIX], pp[SHIFT_MAX_PREFIX];
    unsigned char pclass[SHIFT_MAX_PREFIX + 1];
    int j = 0, L = -1, dr = -1, jend = 0;
    bool prev_diffuse = false, offsets_ok = true;
    double C[3], P = 1, inv_rr = 1, Sf[3], Sp = 1, contrib[3], cS[4][3], wS[4];
    V3<R> Le(0, 0, 0), wr(0, 0, 1), pos_dr(0, 0, 0), rad(0, 0, 0);
    Vertex<R> vr;
    // offset in flight
    int k = 0, oj = 0;
    double Cs[3], Ps = 1, J = 1;

    for (;;) {
        if (phase == SH_FETCH) {
            i = atomicAdd(next_sample, 1);
            if (i >= wp.n_paths) break;
            unsigned long long stream;
            if (wp.xs) { px = wp.xs[i]; py = wp.ys[i]; stream = wp.streams[i]; }
            else {
                long long s = i / wp.npix_pad, p = i % wp.npix_pad;
                int tile = (int)(p >> 5), lane = (int)(p & 31);
                px = (tile % wp.tiles_x) * 8 + (lane & 7); py = wave_row(wp, tile / wp.tiles_x, lane);
                if (px >= wp.width || py >= wp.y1) continue;
                stream = (unsigned long long)(wp.sample0 + s) * (unsigned long long)wp.width * wp.height + (unsigned long long)py * wp.width + px;
            }
            for (int q = 0; q < 4; q++) { cS[q][0] = cS[q][1] = cS[q][2] = 0.0; wS[q] = 1.0; }
            for (int c = 0; c < 3; c++) { C[c] = 1; Sf[c] = 1; contrib[c] = 0; }
            P = 1; inv_rr = 1; Sp = 1; L = -1; dr = -1; j = 0; prev_diffuse = false; offsets_ok = true;
            rad = V3<R>(0, 0, 0); Le = V3<R>(0, 0, 0);
            rng = pcg_init(stream, wp.seed);
            rx = pcg_real<R>(rng); ry = pcg_real<R>(rng);
            sample_primary<R>(sc, (R(px) + rx) / R(sc.width), (R(py) + ry) / R(sc.height), ro, rd);
            pos_dr = ro;
            rtnear = 0.f; rtfar = inf; rany = false; phase = SH_BASE;
        }
        // ---- the one ray-cast site
        Hit h;
        {
            const float of[3] = {(float)ro.x, (float)ro.y, (float)ro.z}, df[3] = {(float)rd.x, (float)rd.y, (float)rd.z};
            n_rays++;
            h = shift_trace<EXACT>(bvh, of, df, rtnear, rtfar, rany, stack);
        }
        bool next_offset = false;
        if (phase == SH_BASE) {
            bool over = true, base_ok = false;
            if (h.shape_id >= 0) {
                Vertex<R> v;
                make_vertex_ms<R, MS>(sc, ro, rd, rtnear, inf, h, R(0), j == 0 ? spread0 : R(0), v);
                const int cls = shift_class(sc, v);
                if (dr < 0 && j > 0 && prev_diffuse && (cls & 1)) { dr = j - 1; vr = v; pos_dr = ro; }
                if (dr < 0 || j <= dr) { if (j <= SHIFT_MAX_PREFIX) pclass[j] = (unsigned char)cls; else offsets_ok = false; }
                if (cls & 2) { Le = emission<R>(sc, v, -rd); L = j; base_ok = true; }
                else if (j + 1 < SHIFT_MAX_VERTS && !(sc.max_depth != -1 && j + 2 > sc.max_depth)) {
                    const R u0 = pcg_real<R>(rng), u1 = pcg_real<R>(rng), w = pcg_real<R>(rng);
                    BsdfSample<R> s;
                    if (bsdf_sample_ms<R, MS>(sc, v, -rd, u0, u1, w, s)) {
                        const V3<R> f = bsdf_eval_ms<R, MS>(sc, v, -rd, s.dir_out);
                        const R p = bsdf_pdf_ms<R, MS>(sc, v, -rd, s.dir_out);
                        if (p > R(0) && (f.x > R(0) || f.y > R(0) || f.z > R(0))) {
                            if (dr < 0) { if (j < SHIFT_MAX_PREFIX) { pu0[j] = u0; pu1[j] = u1; pw[j] = w; pp[j] = p; } else offsets_ok = false; }
                            else if (j == dr + 1) wr = s.dir_out;
                            else { Sf[0] *= (double)f.x; Sf[1] *= (double)f.y; Sf[2] *= (double)f.z; Sp *= (double)p; }
                            C[0] *= (double)f.x; C[1] *= (double)f.y; C[2] *= (double)f.z; P *= (double)p;
                            bool alive = true;
                            if (j + 2 >= sc.rr_depth) {
                                if (!(pcg_real<R>(rng) <= R(SHIFT_RR_SURVIVAL))) alive = false;
                                else inv_rr /= SHIFT_RR_SURVIVAL;
                            }
                            if (alive) {
                                prev_diffuse = (cls & 1) != 0;
                                ro = v.position; rd = s.dir_out; rtnear = (float)sc.eps; j++;
                                over = false;
                            }
                        }
                    }
                }
            }
            if (!over) continue;
            if (!base_ok) {
                if (wp.rec_out) finish_path<R>(wp, px, py, i, rad, contrib, 1.0, cS, wS);
                phase = SH_FETCH; continue;
            }
            for (int c = 0; c < 3; c++) contrib[c] = C[c] * (double)(c == 0 ? Le.x : c == 1 ? Le.y : Le.z) * inv_rr;
            rad = V3<R>((R)(contrib[0] / P), (R)(contrib[1] / P), (R)(contrib[2] / P));
            jend = dr >= 0 ? dr : L;
            k = -1; next_offset = true;
        } else if (phase == SH_OFFSET) {
            bool ok = h.shape_id >= 0, done = false;
            if (ok) {
                Vertex<R> vo;
                make_vertex_ms<R, MS>(sc, ro, rd, rtnear, inf, h, R(0), oj == 0 ? spread0 : R(0), vo);
                if (shift_class(sc, vo) != (int)pclass[oj]) ok = false;
                else if (oj < jend) {                             // replay the base's random numbers at this vertex
                    BsdfSample<R> s;
                    if (!bsdf_sample_ms<R, MS>(sc, vo, -rd, pu0[oj], pu1[oj], pw[oj], s)) ok = false;
                    else {
                        const V3<R> f = bsdf_eval_ms<R, MS>(sc, vo, -rd, s.dir_out);
                        const R p = bsdf_pdf_ms<R, MS>(sc, vo, -rd, s.dir_out);
                        if (!(p > R(0))) ok = false;
                        else {
                            Cs[0] *= (double)f.x; Cs[1] *= (double)f.y; Cs[2] *= (double)f.z; Ps *= (double)p; J *= (double)pp[oj] / (double)p;
                            ro = vo.position; rd = s.dir_out; rtnear = (float)sc.eps; oj++;
                            continue;
                        }
                    }
                } else if (dr < 0) {                              // replayed all the way: the offset's own light vertex
                    const V3<R> e = emission<R>(sc, vo, -rd);
                    Cs[0] *= (double)e.x; Cs[1] *= (double)e.y; Cs[2] *= (double)e.z;
                    done = true;
                } else {                                          // reconnect vo -> b_{dr+1}: everything but the visibility ray
                    const V3<R> dv = vr.position - vo.position;
                    const R d2s = dot(dv, dv);
                    ok = d2s > R(0);
                    if (ok) {
                        const R dist = sqrt(d2s);
                        const V3<R> ds = dv / dist;
                        const V3<R> f = bsdf_eval_ms<R, MS>(sc, vo, -rd, ds);
                        const R p = bsdf_pdf_ms<R, MS>(sc, vo, -rd, ds);
                        const V3<R> bv = vr.position - pos_dr;
                        const R d2b = dot(bv, bv);
                        const R cos_s = fabs(dot(ds, vr.gn)), cos_b = fabs(dot(bv, vr.gn)) / sqrt(d2b);
                        ok = p > R(0) && cos_b > R(0) && d2b > R(0);
                        if (ok) {
                            J *= ((double)cos_s / (double)d2s) / ((double)cos_b / (double)d2b);
                            Cs[0] *= (double)f.x; Cs[1] *= (double)f.y; Cs[2] *= (double)f.z; Ps *= (double)p;
                            if (L == dr + 1) {
                                const V3<R> e = emission<R>(sc, vr, -ds);
                                Cs[0] *= (double)e.x; Cs[1] *= (double)e.y; Cs[2] *= (double)e.z;
                            } else {
                                const V3<R> ft = bsdf_eval_ms<R, MS>(sc, vr, -ds, wr);
                                const R pt = bsdf_pdf_ms<R, MS>(sc, vr, -ds, wr);
                                ok = pt > R(0);
                                Cs[0] *= (double)ft.x * Sf[0] * (double)Le.x; Cs[1] *= (double)ft.y * Sf[1] * (double)Le.y; Cs[2] *= (double)ft.z * Sf[2] * (double)Le.z;
                                Ps *= (double)pt * Sp;
                            }
                        }
                        if (ok) {
                            ro = vo.position; rd = ds; rtnear = (float)sc.eps; rtfar = (float)((R(1) - sc.eps) * dist); rany = true;
                            phase = SH_RECONNECT;
                            continue;
                        }
                    }
                }
            }
            if (ok && done && J > 0.0 && isfinite(J) && isfinite(Ps)) {
                for (int c = 0; c < 3; c++) cS[k][c] = Cs[c] * J * inv_rr;
                wS[k] = P / (P + Ps * J);
            }
            next_offset = true;
        } else {                                                  // SH_RECONNECT: the visibility ray is back
            if (h.shape_id < 0 && J > 0.0 && isfinite(J) && isfinite(Ps)) {
                for (int c = 0; c < 3; c++) cS[k][c] = Cs[c] * J * inv_rr;
                wS[k] = P / (P + Ps * J);
            }
            next_offset = true;
        }
        if (next_offset) {
            for (k++; k < 4; k++) {
                if (!offsets_ok) { k = 4; break; }
                if (!wp.rec_out && (px + ox[k] < 0 || px + ox[k] >= wp.width || py + oy[k] < 0 || py + oy[k] >= wp.height)) continue;
                break;
            }
            if (k < 4) {
                sample_primary<R>(sc, (R(px + ox[k]) + rx) / R(sc.width), (R(py + oy[k]) + ry) / R(sc.height), ro, rd);
                rtnear = 0.f; rtfar = inf; rany = false; oj = 0;
                Cs[0] = Cs[1] = Cs[2] = 1; Ps = 1; J = 1;
                phase = SH_OFFSET;
            } else {
                finish_shift<R>(wp, px, py, i, rad, contrib, P, cS, wS);
                phase = SH_FETCH;
            }
        }
    }
    for (int s = 16; s > 0; s >>= 1) n_rays += __shfl_xor_sync(0xffffffffu, n_rays, s);
    if ((threadIdx.x & 31) == 0 && n_rays) atomicAdd(ray_counter, (unsigned long long)n_rays);
}

template <typename R, int MS>
int run_shift_wave_ms(gdpt_ctx *c, DScene<R> sc, WaveParams wp) {
    const int N = wp.n_paths;
    if (N <= 0) return 0;
    if (int r = c->counters.ensure(64)) return r;
    unsigned long long *cnt = reinterpret_cast<unsigned long long *>(c->counters.as<char>() + 32);
    cudaStream_t st = c->stream;
    cudaEvent_t e0 = c->ev[4], e1 = c->ev[5];
    GD_CUDA(cudaMemsetAsync(cnt, 0, 2 * sizeof(unsigned long long), st));      // rays traced; next sample (k_shift_sm)
    wp.res = nullptr;
    if (c->deterministic && !wp.rec_out && !wp.xs) {      // image renders: per-path result records + ordered reduction (no atomics)
        if (int r = c->res0.ensure((size_t)N * 15 * sizeof(double))) return r;
        wp.res = c->res0.as<double>();
        GD_CUDA(cudaMemsetAsync(wp.res, 0, (size_t)N * 15 * sizeof(double), st));
    }
    GD_CUDA(cudaEventRecord(e0, st));
    // default: the state machine at 8 blocks per SM (64 registers; B200, Sponza 1080p, 8 spp: 87 ms — 5 blocks 100 ms, 3 blocks
    // 117 ms; k_shift 116 ms, 103 ms at 8 blocks per SM).  GDPT_SHIFT_KERNEL=0 runs k_shift (A/B, bit-identity test).
    const bool straight = getenv("GDPT_SHIFT_KERNEL") && atoi(getenv("GDPT_SHIFT_KERNEL")) == 0;
    const int nb = blocks_for(N, TRACE_BLOCK);
    if (straight) k_shift<R, MS, ModeSel<R>::exact><<<nb, TRACE_BLOCK, 0, st>>>(sc, wp, c->bvh, cnt);
    else k_shift_sm<R, MS, ModeSel<R>::exact, 8><<<std::min(nb, SM_COUNT * 8), TRACE_BLOCK, 0, st>>>(sc, wp, c->bvh, cnt, reinterpret_cast<int *>(cnt + 1));
    if (wp.res) { k_shift_reduce<<<blocks_for(wp.npix_pad, 256), 256, 0, st>>>(wp, wp.res, wp.img, wp.cx0, wp.cy0, wp.cx1, wp.cy1); c->stats.kernel_launches++; }
    GD_CUDA(cudaEventRecord(e1, st));
    unsigned long long h = 0;
    GD_CUDA(cudaMemcpyAsync(&h, cnt, sizeof(h), cudaMemcpyDeviceToHost, st));
    GD_CUDA(cudaStreamSynchronize(st));
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    c->stats.ms_trace += ms;                         // trace and shade are one kernel in this mode
    c->stats.kernel_launches++; c->stats.paths += N; c->stats.rays_traced += (long long)h;
    GD_CUDA(cudaGetLastError());
    return 0;
}

template <typename R>
int run_shift_wave(gdpt_ctx *c, DScene<R> sc, WaveParams wp) {
    return c->lambertian_only ? run_shift_wave_ms<R, MS_LAMBERTIAN>(c, sc, wp) : run_shift_wave_ms<R, MS_GENERIC>(c, sc, wp);
}

} // namespace gdpt
