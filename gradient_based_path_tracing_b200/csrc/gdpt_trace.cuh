// Closest-hit traversal of the 4-wide BVH (gdpt_bvh.h) — the replacement for
// rtcIntersect1 (/root/reference/src/intersection.cpp:15-36) and the sphere user-geometry
// callback (/root/reference/src/shapes/sphere.inl:40-106).
//
//  * one thread per ray; the ray is the FLOAT ray Embree would be handed;
//  * "while-while" traversal: all lanes of a warp walk inner nodes until each holds a leaf
//    (or is done), then all test leaves — the two phases never interleave inside a warp;
//  * node = one 128-byte line: six box planes fetched as 3 x 32-byte loads (LDG.E.ENL2.256) + the child ids, 4 slab tests;
//  * children visited front to back; the traversal stack lives in shared memory
//    ([level][thread] -> bank-conflict free) with a small local-memory spill;
//  * hit rule: tnear < t <= tfar, closest t, ties -> smaller shape id, then smaller
//    primitive id (same rule as the oracle's Embree stand-in);
//  * EXACT=true evaluates Moeller-Trumbore in double on the float vertices (bit-comparable
//    with the oracle), EXACT=false in float.
#pragma once
#include "gdpt_scene.cuh"

namespace gdpt {

struct DBvh {
    const float4 *nodes;   // 8 float4 per node
    const float4 *prims;   // 3 float4 per primitive
    const double *spheres; // [n_shapes][4]
    int n_nodes, n_prims;
};

constexpr int TRACE_BLOCK = 128;
constexpr int STACK_SMEM = 20;     // entries per thread in shared memory
constexpr int STACK_TOTAL = 64;    // entries per thread in all (the rest spills to local memory)

template <int DEPTH = STACK_SMEM>
struct TraceStackT {
    int2 *smem;            // [DEPTH][TRACE_BLOCK]
    int2 spill[STACK_TOTAL - DEPTH];
    GD_D void put(int sp, int node, float d) {
        int2 e = make_int2(node, __float_as_int(d));
        if (sp < DEPTH) smem[sp * TRACE_BLOCK + threadIdx.x] = e; else spill[sp - DEPTH] = e;
    }
    GD_D int2 get(int sp) const { return sp < DEPTH ? smem[sp * TRACE_BLOCK + threadIdx.x] : spill[sp - DEPTH]; }
};
using TraceStack = TraceStackT<STACK_SMEM>;

// r: the ray state (o, d, tnear, best = current tfar, hit; EXACT mode also best_d).  EXACT mode decides "closer" in DOUBLE like
// the oracle's ray caster — accepted iff tnear < t <= best_d, an exactly equal depth goes to the smaller (shape, primitive) id —
// so the hit does not depend on the order in which the tree presents the primitives; best (float, rounded from best_d) only
// prunes boxes.  A sphere is tested by the reference's own callback against the FLOAT tfar of the Embree ray (sphere.inl:77-85),
// and its hit distance travels on as that float (intersection.cpp:39-40): mirrored here.
template <bool EXACT, bool FLAT = false, typename RS>
GD_D void test_prim(const DBvh &bvh, int pi, RS &r) {
    const float *o = r.o, *d = r.d; const float tnear = r.tnear; float &best = r.best; Hit &hit = r.hit;
    float4 a = __ldg(bvh.prims + 3 * pi), b = __ldg(bvh.prims + 3 * pi + 1), c = __ldg(bvh.prims + 3 * pi + 2);
    int prim_id = __float_as_int(a.w), shape_id = __float_as_int(b.w), kind = __float_as_int(c.w);
    if (kind == 1) {
        double t = sphere_hit_d(bvh.spheres + 4 * shape_id, o, d, tnear, best);
        if (t >= 0.0) { best = (float)t; r.best_d = (double)best; hit.t = best; hit.u = 0.f; hit.v = 0.f; hit.shape_id = shape_id; hit.prim_id = prim_id; }
        return;
    }
    float t, u, v;
    if (!EXACT && FLAT) {
        // same arithmetic as below without the early exits: one predicate at the end (a zero or NaN determinant
        // turns u, v, t into inf/NaN and fails the comparisons), so a leaf costs one straight run of instructions
        float e1x = b.x - a.x, e1y = b.y - a.y, e1z = b.z - a.z;
        float e2x = c.x - a.x, e2y = c.y - a.y, e2z = c.z - a.z;
        float px = d[1] * e2z - d[2] * e2y, py = d[2] * e2x - d[0] * e2z, pz = d[0] * e2y - d[1] * e2x;
        float det = e1x * px + e1y * py + e1z * pz;
        float inv = 1.0f / det;
        float sx = o[0] - a.x, sy = o[1] - a.y, sz = o[2] - a.z;
        u = (sx * px + sy * py + sz * pz) * inv;
        float qx = sy * e1z - sz * e1y, qy = sz * e1x - sx * e1z, qz = sx * e1y - sy * e1x;
        v = (d[0] * qx + d[1] * qy + d[2] * qz) * inv;
        t = (e2x * qx + e2y * qy + e2z * qz) * inv;
        bool ok = det != 0.f && u >= 0.f && u <= 1.f && v >= 0.f && u + v <= 1.f && t > tnear && t <= best;
        ok = ok && !(t == best && hit.shape_id >= 0 && !(shape_id < hit.shape_id || (shape_id == hit.shape_id && prim_id < hit.prim_id)));
        if (!ok) return;
    } else if (EXACT) {
        double e1x = (double)b.x - a.x, e1y = (double)b.y - a.y, e1z = (double)b.z - a.z;
        double e2x = (double)c.x - a.x, e2y = (double)c.y - a.y, e2z = (double)c.z - a.z;
        double dx = d[0], dy = d[1], dz = d[2];
        double px = dy * e2z - dz * e2y, py = dz * e2x - dx * e2z, pz = dx * e2y - dy * e2x;
        double det = e1x * px + e1y * py + e1z * pz;
        if (det == 0.0 || det != det) return;
        double inv = 1.0 / det;
        double sx = (double)o[0] - a.x, sy = (double)o[1] - a.y, sz = (double)o[2] - a.z;
        double ud = (sx * px + sy * py + sz * pz) * inv;
        if (!(ud >= 0.0 && ud <= 1.0)) return;
        double qx = sy * e1z - sz * e1y, qy = sz * e1x - sx * e1z, qz = sx * e1y - sy * e1x;
        double vd = (dx * qx + dy * qy + dz * qz) * inv;
        if (!(vd >= 0.0 && ud + vd <= 1.0)) return;
        double td = (e2x * qx + e2y * qy + e2z * qz) * inv;
        if (!(td > (double)tnear && td <= r.best_d)) return;
        if (td == r.best_d && hit.shape_id >= 0 && !(shape_id < hit.shape_id || (shape_id == hit.shape_id && prim_id < hit.prim_id))) return;
        r.best_d = td;
        t = (float)td; u = (float)ud; v = (float)vd;
    } else {
        float e1x = b.x - a.x, e1y = b.y - a.y, e1z = b.z - a.z;
        float e2x = c.x - a.x, e2y = c.y - a.y, e2z = c.z - a.z;
        float px = d[1] * e2z - d[2] * e2y, py = d[2] * e2x - d[0] * e2z, pz = d[0] * e2y - d[1] * e2x;
        float det = e1x * px + e1y * py + e1z * pz;
        if (det == 0.f || det != det) return;
        float inv = 1.0f / det;
        float sx = o[0] - a.x, sy = o[1] - a.y, sz = o[2] - a.z;
        u = (sx * px + sy * py + sz * pz) * inv;
        if (!(u >= 0.f && u <= 1.f)) return;
        float qx = sy * e1z - sz * e1y, qy = sz * e1x - sx * e1z, qz = sx * e1y - sy * e1x;
        v = (d[0] * qx + d[1] * qy + d[2] * qz) * inv;
        if (!(v >= 0.f && u + v <= 1.f)) return;
        t = (e2x * qx + e2y * qy + e2z * qz) * inv;
        if (!(t > tnear && t <= best)) return;
        if (t == best && hit.shape_id >= 0 && !(shape_id < hit.shape_id || (shape_id == hit.shape_id && prim_id < hit.prim_id))) return;
    }
    best = t; hit.t = t; hit.u = u; hit.v = v; hit.shape_id = shape_id; hit.prim_id = prim_id;
}

constexpr int TRACE_DONE = (int)0x80000000;   // "no more work for this ray" (never a valid leaf encoding)
constexpr int TRACE_POP = (int)0x80000001;    // "take the next stack entry" (loop form with one shared pop site; never a valid leaf encoding either)

// Per-ray traversal state kept in registers.
struct RayState {
    float o[3], d[3], id[3];   // origin, direction, 1/direction
    float tnear, best;         // best = current tfar
    double best_d;             // EXACT mode: the accepted hit distance in double (test_prim); unused otherwise
    float oid[3], slack;       // FMA slab form: o*id and the absolute rounding slack that keeps it conservative
    Hit hit;
    GD_D void init(const float4 &a, const float4 &b) {
        o[0] = a.x; o[1] = a.y; o[2] = a.z; d[0] = b.x; d[1] = b.y; d[2] = b.z;
        id[0] = 1.0f / b.x; id[1] = 1.0f / b.y; id[2] = 1.0f / b.z;
        tnear = a.w; best = b.w; best_d = (double)b.w;
        oid[0] = o[0] * id[0]; oid[1] = o[1] * id[1]; oid[2] = o[2] * id[2];
        slack = 2.4e-7f * fmaxf(fmaxf(fabsf(oid[0]), fabsf(oid[1])), fabsf(oid[2]));
        hit.t = b.w; hit.u = hit.v = 0.f; hit.shape_id = -1; hit.prim_id = -1;
    }
};

// next entry of the stack that can still contain a closer hit
template <typename ST>
GD_D int trace_pop(const ST &st, int &sp, float best) {
    while (sp > 0) {
        int2 e = st.get(--sp);
        if (__int_as_float(e.y) <= best) return e.x;
    }
    return TRACE_DONE;
}

// Inner node: slab-test the 4 children, descend into the nearest, push the rest far-to-near.
// EAGER = 1 (EXACT_E selects the triangle arithmetic): children that are leaves and pass the slab test are tested right here,
// in the node step, instead of being visited in a trip of their own — the leaf code then runs for every lane whose node has a
// hit leaf child at this trip, not for the two or three lanes of a warp that happen to sit on a leaf.
// 256-bit global load (sm_100: LDG.E.ENL2.256): two of the node's float4 planes per instruction and per L1 wavefront.
GD_D void ld256(const float4 *p, float4 &a, float4 &b) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}
// the same with the L1 eviction priority "evict last": node lines are the re-used data of the walk, rays / hits / primitives are not
GD_D void ld256_keep(const float4 *p, float4 &a, float4 &b) {
    asm volatile("ld.global.nc.L1::evict_last.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}
GD_D int4 ld128_keep(const int4 *p) {
    int4 r;
    asm volatile("ld.global.nc.L1::evict_last.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
// WIDE: the node's six planes are fetched with three 256-bit loads instead of six 128-bit ones; WIDE = 2: with evict-last priority.
template <int SORT = 0, int FMA = 0, bool NOPOP = false, int EAGER = 0, bool EXACT_E = false, int WIDE = 0, typename ST, typename RS>
GD_D int trace_node(const DBvh &bvh, int cur, RS &r, ST &st, int &sp, unsigned *n_prims = nullptr) {
    const float4 *n = bvh.nodes + 8 * (size_t)cur;
    const float INF = __int_as_float(0x7f800000);
    float dist[4]; int cid[4];
    if (FMA == 2) {
        // near/far planes picked by the sign of 1/d at load time (no min/max per axis), one FMA per plane
        const int sx = __float_as_int(r.id[0]) < 0 ? 3 : 0, sy = __float_as_int(r.id[1]) < 0 ? 3 : 0, sz = __float_as_int(r.id[2]) < 0 ? 3 : 0;
        float4 nx = __ldg(n + sx), ny = __ldg(n + 1 + sy), nz = __ldg(n + 2 + sz);
        float4 fx = __ldg(n + 3 - sx), fy = __ldg(n + 4 - sy), fz = __ldg(n + 5 - sz);
        int4 ch = __ldg((const int4 *)(n + 6));
        cid[0] = ch.x; cid[1] = ch.y; cid[2] = ch.z; cid[3] = ch.w;
        const float anx[4] = {nx.x, nx.y, nx.z, nx.w}, any[4] = {ny.x, ny.y, ny.z, ny.w}, anz[4] = {nz.x, nz.y, nz.z, nz.w};
        const float afx[4] = {fx.x, fx.y, fx.z, fx.w}, afy[4] = {fy.x, fy.y, fy.z, fy.w}, afz[4] = {fz.x, fz.y, fz.z, fz.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
            float tmn = fmaxf(fmaxf(fmaf(anx[k], r.id[0], -r.oid[0]), fmaf(any[k], r.id[1], -r.oid[1])), fmaxf(fmaf(anz[k], r.id[2], -r.oid[2]), r.tnear));
            float tmx = fminf(fminf(fmaf(afx[k], r.id[0], -r.oid[0]), fmaf(afy[k], r.id[1], -r.oid[1])), fminf(fmaf(afz[k], r.id[2], -r.oid[2]), r.best));
            tmx = fmaf(tmx, 1.0000005f, r.slack + r.slack);   // conservative: product rounding + the FMA form's cancellation error
            dist[k] = (cid[k] != 0 && tmn <= tmx) ? tmn : INF;
        }
    } else {
    float4 lox, loy, loz, hix, hiy, hiz;
    if (WIDE == 2) { ld256_keep(n, lox, loy); ld256_keep(n + 2, loz, hix); ld256_keep(n + 4, hiy, hiz); }
    else if (WIDE) { ld256(n, lox, loy); ld256(n + 2, loz, hix); ld256(n + 4, hiy, hiz); }
    else { lox = __ldg(n); loy = __ldg(n + 1); loz = __ldg(n + 2); hix = __ldg(n + 3); hiy = __ldg(n + 4); hiz = __ldg(n + 5); }
    int4 ch = WIDE == 2 ? ld128_keep((const int4 *)(n + 6)) : __ldg((const int4 *)(n + 6));
    cid[0] = ch.x; cid[1] = ch.y; cid[2] = ch.z; cid[3] = ch.w;
    const float lx[4] = {lox.x, lox.y, lox.z, lox.w}, ly[4] = {loy.x, loy.y, loy.z, loy.w}, lz[4] = {loz.x, loz.y, loz.z, loz.w};
    const float hx[4] = {hix.x, hix.y, hix.z, hix.w}, hy[4] = {hiy.x, hiy.y, hiy.z, hiy.w}, hz[4] = {hiz.x, hiz.y, hiz.z, hiz.w};
#pragma unroll
    for (int k = 0; k < 4; k++) {
        float t0x, t1x, t0y, t1y, t0z, t1z;
        if (FMA) {   // l*id - o*id in one FMA per plane; the slack below covers the cancellation error
            t0x = fmaf(lx[k], r.id[0], -r.oid[0]); t1x = fmaf(hx[k], r.id[0], -r.oid[0]);
            t0y = fmaf(ly[k], r.id[1], -r.oid[1]); t1y = fmaf(hy[k], r.id[1], -r.oid[1]);
            t0z = fmaf(lz[k], r.id[2], -r.oid[2]); t1z = fmaf(hz[k], r.id[2], -r.oid[2]);
        } else {
            t0x = (lx[k] - r.o[0]) * r.id[0]; t1x = (hx[k] - r.o[0]) * r.id[0];
            t0y = (ly[k] - r.o[1]) * r.id[1]; t1y = (hy[k] - r.o[1]) * r.id[1];
            t0z = (lz[k] - r.o[2]) * r.id[2]; t1z = (hz[k] - r.o[2]) * r.id[2];
        }
        float tmn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fmaxf(fminf(t0z, t1z), r.tnear));
        float tmx = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fminf(fmaxf(t0z, t1z), r.best));
        tmx *= 1.0000005f;   // conservative slab (rounding of the three products)
        if (FMA) tmx += r.slack + r.slack;
        dist[k] = (cid[k] != 0 && tmn <= tmx) ? tmn : INF;
    }
    }
    if (EAGER) {
#pragma unroll
        for (int k = 0; k < 4; k++) {
            if (cid[k] < 0 && dist[k] < INF) {
                if (dist[k] <= r.best) {
                    const int enc = ~cid[k], first = enc >> 3, cnt = enc & 7;
                    if (n_prims) *n_prims += cnt;
                    for (int q = 0; q < cnt; q++) test_prim<EXACT_E, false>(bvh, first + q, r);
                }
                dist[k] = INF;
            }
        }
#pragma unroll
        for (int k = 0; k < 4; k++) if (dist[k] > r.best) dist[k] = INF;     // a hit found above may have closed the other children
    }
#define GD_CSWAP(i, j) if (dist[j] < dist[i]) { float td = dist[i]; dist[i] = dist[j]; dist[j] = td; int tc = cid[i]; cid[i] = cid[j]; cid[j] = tc; }
    if (SORT == 0) {   // sort the 4 candidates by distance (network), nearest first
        GD_CSWAP(0, 1) GD_CSWAP(2, 3) GD_CSWAP(0, 2) GD_CSWAP(1, 3) GD_CSWAP(1, 2)
    } else {           // nearest first only; the others are culled by distance when popped
        GD_CSWAP(0, 1) GD_CSWAP(0, 2) GD_CSWAP(0, 3)
    }
#undef GD_CSWAP
    if (dist[3] < INF) st.put(sp++, cid[3], dist[3]);
    if (dist[2] < INF) st.put(sp++, cid[2], dist[2]);
    if (dist[1] < INF) st.put(sp++, cid[1], dist[1]);
    if (dist[0] < INF) return cid[0];
    if (NOPOP) return TRACE_POP;
    return trace_pop(st, sp, r.best);
}

// Leaf (cur < 0, != TRACE_DONE): test its primitives, then pop.
template <bool EXACT, bool FLAT = false, typename ST>
GD_D int trace_leaf(const DBvh &bvh, int cur, RayState &r, const ST &st, int &sp, unsigned *n_prims) {
    int enc = ~cur;
    int first = enc >> 3, cnt = enc & 7;
    if (n_prims) *n_prims += cnt;
    for (int i = 0; i < cnt; i++) test_prim<EXACT, FLAT>(bvh, first + i, r);
    return trace_pop(st, sp, r.best);
}

// One ray, one thread (validation entry points).  smem_stack: TRACE_BLOCK * STACK_SMEM int2 owned by the block.
template <bool EXACT, bool COUNT = false>
GD_D Hit trace_closest(const DBvh &bvh, const float o[3], const float d[3], float tnear, float tfar, int2 *smem_stack,
                       unsigned *n_nodes = nullptr, unsigned *n_prims = nullptr) {
    RayState r; r.init(make_float4(o[0], o[1], o[2], tnear), make_float4(d[0], d[1], d[2], tfar));
    if (bvh.n_nodes == 0) return r.hit;
    TraceStack st; st.smem = smem_stack;
    int sp = 0, cur = 0;
    while (cur != TRACE_DONE) {
        while (cur >= 0) { if (COUNT) (*n_nodes)++; cur = trace_node<0, 0, false, 0, false, 1>(bvh, cur, r, st, sp); }
        if (cur != TRACE_DONE) cur = trace_leaf<EXACT>(bvh, cur, r, st, sp, COUNT ? n_prims : nullptr);
    }
    return r.hit;
}

// Occlusion query for one ray (rtcOccluded1): stops at the first accepted hit.
template <bool EXACT>
GD_D bool trace_occluded(const DBvh &bvh, const float o[3], const float d[3], float tnear, float tfar, int2 *smem_stack) {
    RayState r; r.init(make_float4(o[0], o[1], o[2], tnear), make_float4(d[0], d[1], d[2], tfar));
    if (bvh.n_nodes == 0) return false;
    TraceStack st; st.smem = smem_stack;
    int sp = 0, cur = 0;
    while (cur != TRACE_DONE) {
        while (cur >= 0) cur = trace_node<0, 0, false, 0, false, 1>(bvh, cur, r, st, sp);
        if (cur != TRACE_DONE) cur = trace_leaf<EXACT>(bvh, cur, r, st, sp, nullptr);
        if (r.hit.shape_id >= 0) return true;
    }
    return false;
}

} // namespace gdpt
