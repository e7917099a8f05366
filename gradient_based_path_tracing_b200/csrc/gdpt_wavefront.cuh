// Wavefront gradient-domain path tracer: the device side of gradient_path_render
// (/root/reference/src/render.cpp:257-370) and grad_path_tracing
// (/root/reference/src/path_tracing.h:354-1050), "reference" offset semantics
// (the as-shipped behaviour, UB resolved as "every offset still
// enabled when the base path enters its second bounce is dropped with contribS=0, w=1").
//
// One wave = N pixel-samples.  Kernels (all templated on the arithmetic type R):
//   k_generate   rng_x, rng_y and the 5 primary rays (base + X0,Y0,X1,Y1)   path_tracing.h:360-403
//   k_trace      closest hit for a ray queue (gdpt_trace.cuh)                intersection.cpp:7-36
//   k_primary    5 primary vertices, offset enabling, first BSDF sample      path_tracing.h:375-564, 774-959
//   k_bounce     one path-tracing iteration given the hit of the sampled ray path_tracing.h:746-1010
//                + the next BSDF sample; finished paths are accumulated
//                (render.cpp:311-327), survivors are compacted with a warp ballot + one
//                atomic per warp into the next queue.
// Per-path state is a 32-byte-aligned record (AoS: each thread streams whole sectors).
#pragma once
#include "gdpt_bsdf.cuh"
#include "gdpt_trace.cuh"

namespace gdpt {

struct alignas(16) RayF { float ox, oy, oz, tnear, dx, dy, dz, tfar; };   // 32 B, two float4

template <typename R> struct alignas(32) PathRec {
    V3<R> org, dir;          // the ray being traced: org = current vertex position, dir = sampled direction
    V3<R> f;                 // BSDF*cos of the current vertex towards dir
    R p2;                    // solid-angle pdf of dir
    V3<R> thr, rad;          // current_path_throughput, accumulated radiance
    R eta_scale;
    double contrib[3], prob; // whole-walk products: always double (they leave float range quickly)
    Pcg rng;
    int px, py, id, nv;      // pixel, path id (record mode), num_vertices of the iteration in flight
};

template <typename R> struct alignas(32) OffRec {   // the four pixel-shifted offsets, alive during bounce 1 only
    V3<R> contrib[4];        // 1 or the primary-hit emission (path_tracing.h:497-508)
    R p2s[4];                // solid-angle pdf of the replayed sample at the offset vertex
    int enabled;             // bit k: offset k still enabled (check_path_*)
    int sampled;             // bit k: p2s[k] is valid (sample succeeded with pdf > 0)
};

template <typename R> struct alignas(16) GenRec { R rng_x, rng_y; Pcg rng; int px, py, id; };

struct WaveParams {
    int width, height;       // full image
    int y0, y1;              // row band rendered by this context
    int tiles_x;             // ceil(width / 8)
    int n_paths;             // padded path slots in the wave
    long long npix_pad;      // padded pixels per sample pass (tiles * 32)
    int sample0, n_samples;  // first sample index of the wave, samples in the wave
    int spp_total;           // divisor of the estimator (render.cpp:313-317)
    unsigned long long seed;
    // record mode (parity tests): explicit pixel/stream lists, results to rec_out[id*23]
    const int *xs, *ys; const unsigned long long *streams; double *rec_out;
    // image mode: five accumulation images, W*H*3 doubles each, row-major, y down
    double *img, *cx0, *cy0, *cx1, *cy1;
};

// Records move between registers and HBM as 16-byte vectors (the compiler's own struct copy
// degenerates to byte accesses for these aggregates).
template <typename T> GD_D void load_rec(T &dst, const T *src) {
    static_assert(sizeof(T) % 16 == 0 && alignof(T) >= 16, "record must be a multiple of 16 bytes");
    const int4 *s = reinterpret_cast<const int4 *>(src);
    int4 *d = reinterpret_cast<int4 *>(&dst);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); i++) d[i] = s[i];
}
template <typename T> GD_D void store_rec(T *dst, const T &src) {
    static_assert(sizeof(T) % 16 == 0 && alignof(T) >= 16, "record must be a multiple of 16 bytes");
    int4 *d = reinterpret_cast<int4 *>(dst);
    const int4 *s = reinterpret_cast<const int4 *>(&src);
#pragma unroll
    for (int i = 0; i < (int)(sizeof(T) / 16); i++) d[i] = s[i];
}

GD_D int warp_slot(bool alive, int *counter) {
    unsigned mask = __ballot_sync(0xffffffffu, alive);
    int lane = threadIdx.x & 31;
    int base = 0;
    if (mask) {
        int leader = __ffs(mask) - 1;
        if (lane == leader) base = atomicAdd(counter, __popc(mask));
        base = __shfl_sync(0xffffffffu, base, leader);
    }
    return base + __popc(mask & ((1u << lane) - 1u));
}

// ---------------------------------------------------------------- generate
template <typename R>
__global__ void __launch_bounds__(128) k_generate(DScene<R> sc, WaveParams wp, GenRec<R> *gen, RayF *rays) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= wp.n_paths) return;
    GenRec<R> g; g.id = i; g.px = -1; g.py = -1;
    unsigned long long stream = 0; bool active = false;
    if (wp.xs) {
        g.px = wp.xs[i]; g.py = wp.ys[i]; stream = wp.streams[i]; active = true;
    } else {
        long long s = i / wp.npix_pad, p = i % wp.npix_pad;
        int tile = (int)(p >> 5), lane = (int)(p & 31);
        int x = (tile % wp.tiles_x) * 8 + (lane & 7), y = wp.y0 + (tile / wp.tiles_x) * 4 + (lane >> 3);
        if (x < wp.width && y < wp.y1) {
            g.px = x; g.py = y; active = true;
            stream = (unsigned long long)(wp.sample0 + s) * (unsigned long long)wp.width * wp.height + (unsigned long long)y * wp.width + x;
        }
    }
    RayF r; r.ox = r.oy = r.oz = 0.f; r.dx = r.dy = r.dz = 0.f; r.tnear = 0.f; r.tfar = -1.f;   // tfar < tnear: never hits
    if (!active) {
        store_rec(gen + i, g);
        for (int k = 0; k < 5; k++) store_rec(rays + (size_t)k * wp.n_paths + i, r);
        return;
    }
    g.rng = pcg_init(stream, wp.seed);
    g.rng_x = pcg_real<R>(g.rng); g.rng_y = pcg_real<R>(g.rng);
    store_rec(gen + i, g);
    const int ox[5] = {0, -1, 0, 1, 0}, oy[5] = {0, 0, 1, 0, -1};   // base, X0=(x-1,y), Y0=(x,y+1), X1=(x+1,y), Y1=(x,y-1)
#pragma unroll
    for (int k = 0; k < 5; k++) {
        V3<R> o, d;
        sample_primary<R>(sc, (R(g.px + ox[k]) + g.rng_x) / R(sc.width), (R(g.py + oy[k]) + g.rng_y) / R(sc.height), o, d);
        r.ox = (float)o.x; r.oy = (float)o.y; r.oz = (float)o.z; r.dx = (float)d.x; r.dy = (float)d.y; r.dz = (float)d.z;
        r.tnear = 0.f; r.tfar = __int_as_float(0x7f800000);
        store_rec(rays + (size_t)k * wp.n_paths + i, r);
    }
}

// ---------------------------------------------------------------- trace
// Persistent blocks: each block owns a contiguous chunk of the ray queue and its lanes pull
// the next ray from a shared-memory cursor the moment their current ray finishes, so a warp
// is not held hostage by its longest ray.  COUNT=true additionally sums visited nodes /
// tested primitives into counters[0..1] (used once by bench.py to derive the algorithmic
// bytes per ray; never in timed runs).
template <bool EXACT, bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK) k_trace(DBvh bvh, const RayF *rays, Hit *hits, const int *count_ptr, int count_max,
                                                       unsigned long long *counters) {
    __shared__ int2 stack[STACK_SMEM * TRACE_BLOCK];
    __shared__ int cursor;
    const unsigned FULL = 0xffffffffu;
    const int n = count_ptr ? min(*count_ptr, count_max) : count_max;
    const long long per = (((long long)n + gridDim.x - 1) / gridDim.x + 31) & ~31LL;
    const int begin = (int)min((long long)n, blockIdx.x * per), end = (int)min((long long)n, begin + per);
    if (threadIdx.x == 0) cursor = begin;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    TraceStack st; st.smem = stack;
    RayState r;
    int ray = -1, cur = TRACE_DONE, sp = 0;
    bool exhausted = false;
    unsigned cn = 0, cp = 0;
    unsigned long long tn = 0, tp = 0;
    while (true) {
        const bool need = cur == TRACE_DONE && !exhausted;
        const unsigned m = __ballot_sync(FULL, need);
        if (m) {
            int base = 0;
            const int leader = __ffs(m) - 1;
            if (lane == leader) base = atomicAdd(&cursor, __popc(m));
            base = __shfl_sync(FULL, base, leader);
            if (need) {
                if (ray >= 0) { hits[ray] = r.hit; if (COUNT) { tn += cn; tp += cp; cn = cp = 0; } }
                ray = base + __popc(m & ((1u << lane) - 1u));
                if (ray < end) {
                    const float4 *rp = (const float4 *)(rays + ray);
                    float4 a = __ldg(rp), b = __ldg(rp + 1);
                    r.init(a, b);
                    sp = 0;
                    cur = (b.w < a.w || bvh.n_nodes == 0) ? TRACE_DONE : 0;   // tfar < tnear: never hits
                } else { ray = -1; exhausted = true; }
            }
        }
        if (__all_sync(FULL, cur == TRACE_DONE && exhausted)) break;
        while (cur >= 0) { if (COUNT) cn++; cur = trace_node(bvh, cur, r, st, sp); }
        if (cur != TRACE_DONE) cur = trace_leaf<EXACT>(bvh, cur, r, st, sp, COUNT ? &cp : nullptr);
    }
    if (COUNT) {
        for (int s = 16; s > 0; s >>= 1) { tn += __shfl_down_sync(FULL, tn, s); tp += __shfl_down_sync(FULL, tp, s); }
        if (lane == 0) { atomicAdd(counters, tn); atomicAdd(counters + 1, tp); }
    }
}

} // namespace gdpt
