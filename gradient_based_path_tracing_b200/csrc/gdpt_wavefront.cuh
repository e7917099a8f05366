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
    int strip_count, strip_phase;   // of the 4-row strips of [y0, y1) this context renders those with index = phase (mod count); count <= 1: all
    int n_paths;             // padded path slots in the wave
    long long npix_pad;      // padded pixels per sample pass (tiles * 32)
    int sample0, n_samples;  // first sample index of the wave, samples in the wave
    int spp_total;           // divisor of the estimator (render.cpp:313-317)
    unsigned long long seed;
    // record mode (parity tests): explicit pixel/stream lists, results to rec_out[id*23]
    const int *xs, *ys; const unsigned long long *streams; double *rec_out;
    // image mode: five accumulation images, W*H*3 doubles each, row-major, y down
    double *img, *cx0, *cy0, *cx1, *cy1;
    // deterministic accumulation (gradpath, reference offsets): every finished path writes its 15 contributions (radiance3,
    // then X0, Y0, X1, Y1 gradients x 3) to res[id * 15 ..] instead of adding them to the images with atomics; k_wave_reduce then
    // adds them pixel by pixel in sample order.  nullptr: atomicAdd(double) straight into the images.
    double *res;
    // ray sorting (set by run_wave when enabled): per-slot key + histogram, origin grid over the scene box
    int *sort_keys, *sort_hist;
    float sort_lo[3], sort_scale[3];
    int sort_cell_bits, sort_dir_bits;
};

// row of lane `lane` of the 8x4 pixel tile in tile row `tile_row` of this context's share of the band
GD_D int wave_row(const WaveParams &wp, int tile_row, int lane) {
    const int strip = wp.strip_count > 1 ? wp.strip_phase + tile_row * wp.strip_count : tile_row;
    return wp.y0 + strip * 4 + (lane >> 3);
}

// ---------------------------------------------------------------- ray sorting
// Bounce rays are incoherent; before they are traced the queue is re-ordered by
// key = (coarse origin cell, Morton) : (direction, 8x8 octahedral bins) : (fine cell bit per axis)
// with a counting sort: the shade kernel that emits a ray also bumps the key's histogram
// bin, k_sort_scan turns the histogram into start offsets, k_sort_scatter writes the
// permutation `order` that k_trace follows.  Rays themselves never move.
constexpr int SORT_CELL_BITS = 5;                       // default origin grid: 32 cells per axis over the scene box
constexpr int SORT_DIR_BITS = 3;                        // default direction grid: 8 x 8 octahedral bins
constexpr int SORT_CHUNK = 32768;                       // histogram bins scanned by one block
constexpr int SORT_MAX_KEY_BITS = 25;                   // 3*cell + 2*dir may not exceed this (1024 chunks)
inline int sort_bins(int cell_bits, int dir_bits) { return 1 << (3 * cell_bits + 2 * dir_bits); }

GD_D unsigned sort_spread(unsigned x, int bits) {   // abcd -> a00b00c00d (up to 10 bits)
    x &= (1u << bits) - 1u;
    x = (x | (x << 16)) & 0x030000ffu; x = (x | (x << 8)) & 0x0300f00fu; x = (x | (x << 4)) & 0x030c30c3u; x = (x | (x << 2)) & 0x09249249u;
    return x;
}
// key = [coarse cell Morton : direction bin : lowest cell bit per axis] — a warp of the sorted queue shares a
// 2^(CELL_BITS-1) cell and a ~22 degree direction bin and differs in the fine cell
GD_D void sort_emit(const WaveParams &wp, int slot, const RayF &r) {
    if (!wp.sort_hist) return;
    const int cb = wp.sort_cell_bits, db = wp.sort_dir_bits;
    const int top = (1 << cb) - 1;
    int cx = min(top, max(0, (int)((r.ox - wp.sort_lo[0]) * wp.sort_scale[0])));
    int cy = min(top, max(0, (int)((r.oy - wp.sort_lo[1]) * wp.sort_scale[1])));
    int cz = min(top, max(0, (int)((r.oz - wp.sort_lo[2]) * wp.sort_scale[2])));
    // octahedral map of the direction to [-1,1]^2
    float inv = 1.0f / (fabsf(r.dx) + fabsf(r.dy) + fabsf(r.dz));
    float u = r.dx * inv, v = r.dy * inv;
    if (r.dz < 0.f) { float tu = (1.f - fabsf(v)) * (u >= 0.f ? 1.f : -1.f), tv = (1.f - fabsf(u)) * (v >= 0.f ? 1.f : -1.f); u = tu; v = tv; }
    const int dtop = (1 << db) - 1;
    int du = min(dtop, max(0, (int)((u * 0.5f + 0.5f) * (float)(1 << db)))), dv = min(dtop, max(0, (int)((v * 0.5f + 0.5f) * (float)(1 << db))));
    unsigned coarse = sort_spread(cx >> 1, cb - 1) | (sort_spread(cy >> 1, cb - 1) << 1) | (sort_spread(cz >> 1, cb - 1) << 2);
    unsigned key = (((coarse << (2 * db)) | (unsigned)(du << db) | (unsigned)dv) << 3) | (unsigned)((cx & 1) | ((cy & 1) << 1) | ((cz & 1) << 2));
    wp.sort_keys[slot] = (int)key;
    atomicAdd(wp.sort_hist + key, 1);
}
// exclusive scan of the histogram: every block scans one SORT_CHUNK-bin chunk in place and leaves its total in
// chunk_sum; k_sort_chunks then scans the totals; k_sort_scatter adds the two
static __global__ void __launch_bounds__(1024) k_sort_scan(int *hist_all, int *chunk_sum) {
    constexpr int PER = SORT_CHUNK / 1024;
    int *hist = hist_all + (size_t)blockIdx.x * SORT_CHUNK;
    __shared__ int warp_sum[32];
    int v[PER], sum = 0;
    int4 *h4 = reinterpret_cast<int4 *>(hist + threadIdx.x * PER);
#pragma unroll
    for (int i = 0; i < PER / 4; i++) { int4 q = h4[i]; v[4 * i] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w; }
#pragma unroll
    for (int i = 0; i < PER; i++) { int t = v[i]; v[i] = sum; sum += t; }
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, incl = sum;
    for (int s = 1; s < 32; s <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, s); if (lane >= s) incl += t; }
    if (lane == 31) warp_sum[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int ws = warp_sum[lane], wi = ws;
        for (int s = 1; s < 32; s <<= 1) { int t = __shfl_up_sync(0xffffffffu, wi, s); if (lane >= s) wi += t; }
        warp_sum[lane] = wi - ws;
        if (lane == 31) chunk_sum[blockIdx.x] = wi;
    }
    __syncthreads();
    int base = warp_sum[wid] + incl - sum;
#pragma unroll
    for (int i = 0; i < PER / 4; i++) h4[i] = make_int4(v[4 * i] + base, v[4 * i + 1] + base, v[4 * i + 2] + base, v[4 * i + 3] + base);
}
static __global__ void __launch_bounds__(1024) k_sort_chunks(int *chunk_sum, int n_chunks) {   // exclusive scan of <= 1024 chunk totals, one block
    __shared__ int warp_sum[32];
    int t = threadIdx.x < n_chunks ? chunk_sum[threadIdx.x] : 0;
    int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, incl = t;
    for (int s = 1; s < 32; s <<= 1) { int q = __shfl_up_sync(0xffffffffu, incl, s); if (lane >= s) incl += q; }
    if (lane == 31) warp_sum[wid] = incl;
    __syncthreads();
    if (wid == 0) {
        int ws = warp_sum[lane], wi = ws;
        for (int s = 1; s < 32; s <<= 1) { int q = __shfl_up_sync(0xffffffffu, wi, s); if (lane >= s) wi += q; }
        warp_sum[lane] = wi - ws;
    }
    __syncthreads();
    if (threadIdx.x < n_chunks) chunk_sum[threadIdx.x] = warp_sum[wid] + incl - t;
}
// n_ptr (optional): the live count on the device, for launches sized by an upper bound (no host read-back)
static __global__ void __launch_bounds__(256) k_sort_scatter(const int *keys, int *offsets, const int *chunk_off, int n, int *order, const int *n_ptr = nullptr) {
    if (n_ptr) n = min(n, *n_ptr);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        int k = keys[i]; order[atomicAdd(offsets + k, 1) + __ldg(chunk_off + k / SORT_CHUNK)] = i;
    }
}

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

// Queue records (PathRec, OffRec, GenRec) live in HBM as planes of 16-byte vectors: vector k of record i sits at
// base[k*n + i] (n = slots in the wave), so a warp moving its 32 records issues fully coalesced 512-byte accesses
// instead of 32 strided ones.  Rays stay AoS (32 B): the sorted trace launch gathers them one by one.
template <typename T> GD_D void load_soa(T &dst, const T *base, int i, int n) {
    static_assert(sizeof(T) % 16 == 0 && alignof(T) >= 16, "record must be a multiple of 16 bytes");
    const int4 *s = reinterpret_cast<const int4 *>(base) + i;
    int4 *d = reinterpret_cast<int4 *>(&dst);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(T) / 16); k++) d[k] = s[(size_t)k * n];
}
template <typename T> GD_D void store_soa(T *base, int i, int n, const T &src) {
    static_assert(sizeof(T) % 16 == 0 && alignof(T) >= 16, "record must be a multiple of 16 bytes");
    int4 *d = reinterpret_cast<int4 *>(base) + i;
    const int4 *s = reinterpret_cast<const int4 *>(&src);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(T) / 16); k++) d[(size_t)k * n] = s[k];
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
        int x = (tile % wp.tiles_x) * 8 + (lane & 7), y = wave_row(wp, tile / wp.tiles_x, lane);
        if (x < wp.width && y < wp.y1) {
            g.px = x; g.py = y; active = true;
            stream = (unsigned long long)(wp.sample0 + s) * (unsigned long long)wp.width * wp.height + (unsigned long long)y * wp.width + x;
        }
    }
    RayF r; r.ox = r.oy = r.oz = 0.f; r.dx = r.dy = r.dz = 0.f; r.tnear = 0.f; r.tfar = -1.f;   // tfar < tnear: never hits
    if (!active) {
        store_soa(gen, i, wp.n_paths, g);
        for (int k = 0; k < 5; k++) store_rec(rays + (size_t)k * wp.n_paths + i, r);
        return;
    }
    g.rng = pcg_init(stream, wp.seed);
    g.rng_x = pcg_real<R>(g.rng); g.rng_y = pcg_real<R>(g.rng);
    store_soa(gen, i, wp.n_paths, g);
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
// One ray per thread, grid-stride, independent lanes.  (Persistent-warp variants with dynamic
// ray fetch — "while-while" and a re-joined node-step/leaf-step loop — were measured 35-75%
// slower on B200: independent thread scheduling hides the node-fetch latency of diverged
// lanes better than forced reconvergence does; DESIGN.md, "what was tried".)
// Tunables V = SORT + 10*FMA + 100*STACKSEL + 1000*OCC (GDPT_TRACE_VARIANT for sweeps):
//   SORT 0 children fully sorted by distance / 1 nearest first only
//   FMA  0 (l-o)*id slabs / 1 fma(l, id, -o*id) with an absolute slack / 2 the same with near/far planes
//        selected by the ray octant at load time (6 instead of 12 min/max per child)
//   STACKSEL 0: 20, 1: 16, 2: 12, 3: 8 shared-memory stack entries per thread (rest spills)
//   OCC  0: launch bound 8 blocks/SM (64 regs), 1: 7 (72), 2: 9 (56), 3: 10 (48)
//   LOOP (x10000) bit 0: a node step that lands on a leaf tests it in the same trip; bit 1: triangle test without early exits;
//        bit 2: one pop site per trip (node misses and finished leaves pop together at the top of the next trip);
//        bit 3 (with bit 2): leaf children are tested inside their parent's node step (no leaf trips);
//        bit 4: streaming (evict-first) loads of rays / permutation and stores of hits;
//        bit 5 (with bit 2): the node planes are fetched with three 256-bit loads (LDG.E.ENL2.256) instead of six 128-bit ones;
//        bit 6 (with bits 2, 5): node loads carry the L1 eviction priority evict-last
// COUNT=true additionally sums visited nodes / tested primitives into counters[0..1]
// (bench.py derives the algorithmic bytes per ray from them; never in timed runs).
constexpr int trace_variant_depth(int v) { return ((v / 100) % 10) == 0 ? 20 : ((v / 100) % 10) == 1 ? 16 : ((v / 100) % 10) == 2 ? 12 : 8; }
constexpr int trace_variant_blocks(int v) { return ((v / 1000) % 10) == 0 ? 8 : ((v / 1000) % 10) == 1 ? 7 : ((v / 1000) % 10) == 2 ? 9 : 10; }

// ANY: occlusion query (rtcOccluded1, intersection.cpp:66-85): the walk stops at the first accepted hit.
template <bool EXACT, bool COUNT, int V, bool ANY = false>
__global__ void __launch_bounds__(TRACE_BLOCK, trace_variant_blocks(V)) k_trace(DBvh bvh, const RayF *rays, HitRec *hits, const int *order, int count_max,
                                                                               unsigned long long *counters, const int *count_ptr = nullptr,
                                                                               unsigned long long *rays_total = nullptr) {
    constexpr int SORT = V % 10, FMA = (V / 10) % 10, DEPTH = trace_variant_depth(V), LOOP = V / 10000;
    __shared__ int2 stack[DEPTH * TRACE_BLOCK];
    const int n = count_ptr ? min(*count_ptr, count_max) : count_max;   // count_ptr: live count on the device (launch sized by an upper bound)
    if (rays_total && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(rays_total, (unsigned long long)n);
    TraceStackT<DEPTH> st; st.smem = stack;
    unsigned long long tn = 0, tp = 0;
    for (int j = blockIdx.x * TRACE_BLOCK + threadIdx.x; j < n; j += gridDim.x * TRACE_BLOCK) {
        const int i = order ? ((LOOP & 16) ? __ldcs(order + j) : __ldg(order + j)) : j;   // sorted launch: follow the permutation, rays stay where they are
        const float4 *rp = (const float4 *)(rays + i);
        // LOOP bit 4: the ray / permutation / hit streams are touched once — streaming loads and stores keep them from
        // displacing the tree in L1 / L2
        float4 a = (LOOP & 16) ? __ldcs(rp) : __ldg(rp), b = (LOOP & 16) ? __ldcs(rp + 1) : __ldg(rp + 1);
        RayState r; r.init(a, b);
        unsigned cn = 0, cp = 0;
        if (!(b.w < a.w) && bvh.n_nodes != 0) {   // tfar < tnear: never hits
            int sp = 0, cur = 0;
            if (LOOP & 4) {   // one pop site at the top of the trip, shared by the lanes coming from a node miss and from a leaf
                while (true) {
                    if (cur == TRACE_POP) { cur = trace_pop(st, sp, r.best); if (cur == TRACE_DONE) break; }
                    if (cur >= 0) {
                        if (COUNT) cn++;
                        cur = trace_node<SORT, FMA, true, (LOOP & 8) ? 1 : 0, EXACT, (LOOP & 64) ? 2 : (LOOP & 32) ? 1 : 0>(bvh, cur, r, st, sp, COUNT ? &cp : nullptr);
                        if ((LOOP & 8) && ANY && r.hit.shape_id >= 0) break;
                    } else {
                        int enc = ~cur, first = enc >> 3, cnt = enc & 7;
                        if (COUNT) cp += cnt;
                        for (int q = 0; q < cnt; q++) test_prim<EXACT, false>(bvh, first + q, r);
                        if (ANY && r.hit.shape_id >= 0) break;
                        cur = TRACE_POP;
                    }
                }
            } else
            while (cur != TRACE_DONE) {
                if (cur >= 0) { if (COUNT) cn++; cur = trace_node<SORT, FMA>(bvh, cur, r, st, sp); if ((LOOP & 1) == 0) continue; }
                if (cur < 0 && cur != TRACE_DONE) cur = trace_leaf<EXACT, (LOOP & 2) != 0>(bvh, cur, r, st, sp, COUNT ? &cp : nullptr);
                if (ANY && r.hit.shape_id >= 0) break;
            }
        }
        if (LOOP & 16) {
            int4 *d = reinterpret_cast<int4 *>(hits + i);
            __stcs(d, make_int4(__float_as_int(r.hit.t), __float_as_int(r.hit.u), __float_as_int(r.hit.v), r.hit.shape_id));
            __stcs(d + 1, make_int4(r.hit.prim_id, 0, 0, 0));
        } else store_hit(hits + i, r.hit);
        if (COUNT) { tn += cn; tp += cp; }
    }
    if (COUNT) {
        for (int s = 16; s > 0; s >>= 1) { tn += __shfl_down_sync(0xffffffffu, tn, s); tp += __shfl_down_sync(0xffffffffu, tp, s); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(counters, tn); atomicAdd(counters + 1, tp); }
    }
}

} // namespace gdpt
