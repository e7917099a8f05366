// Host-side context behind the C ABI: device copies of the flattened scene in both
// arithmetic modes, mip chains, the BVH, wavefront queues and Poisson work buffers.
#pragma once
#include <cstdlib>
#include <chrono>
#include <atomic>
#include <thread>
#include <cstdio>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "../../include/gdpt.h"
#include "gdpt_bvh.h"
#include "gdpt_trace.cuh"

namespace gdpt {

inline std::string &last_error() { static thread_local std::string e; return e; }
inline int fail(int code, const char *fmt, ...) {
    char buf[512]; va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
    last_error() = buf; return code;
}
#define GD_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(GDPT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

// Device buffers come from the device's default stream-ordered memory pool with the release threshold lifted, so a buffer
// released by one context is handed to the next without going back to the driver: creating and destroying a context per
// frame (what a render()-per-call host does, and what bench.py's end-to-end leg times) no longer pays cudaMalloc /
// cudaFree of several GB of queues each time (cudaFree alone was 20-460 ms per gdpt_destroy).  Allocation and release are
// issued on the legacy stream and completed before they return, so every stream may use the memory afterwards.
inline int setup_pool(int device) {
    static bool done[64] = {false};
    if (device < 0 || device >= 64 || done[device]) return 0;
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) != cudaSuccess) { cudaGetLastError(); return 0; }
    unsigned long long keep = ~0ull;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    cudaGetLastError();
    done[device] = true;
    return 0;
}

struct DevBuf {
    void *p = nullptr; size_t bytes = 0;
    int ensure(size_t n) {
        if (n <= bytes) return 0;
        release();
        cudaError_t e = cudaMallocAsync(&p, n, (cudaStream_t)0);
        if (e == cudaSuccess) e = cudaStreamSynchronize((cudaStream_t)0);
        if (e != cudaSuccess) { p = nullptr; return fail(GDPT_ERR_CUDA, "cudaMallocAsync(%zu) failed: %s", n, cudaGetErrorString(e)); }
        bytes = n; return 0;
    }
    void release() {
        if (p) {
            cudaDeviceSynchronize();                  // work on any stream may still use the buffer
            cudaFreeAsync(p, (cudaStream_t)0);
            cudaStreamSynchronize((cudaStream_t)0);
        }
        p = nullptr; bytes = 0;
    }
    template <typename T> T *as() const { return (T *)p; }
};

template <typename T> int upload(DevBuf &b, const std::vector<T> &v) {
    if (int r = b.ensure(std::max<size_t>(v.size() * sizeof(T), 16))) return r;
    if (!v.empty()) {
        GD_CUDA(cudaMemcpy(b.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
        GD_CUDA(cudaStreamSynchronize((cudaStream_t)0));   // pageable source: the call may return before the DMA lands, and the readers run on non-blocking streams
    }
    return 0;
}

struct DctPlanBufs {                  // tables of the FFT form of the DCT-I for one line length (gdpt_poisson_fft.cuh)
    int N = 0, M = 0, P = 0, logP = 0;
    DevBuf tw, chirp, bhat, wpost;
    void release() { tw.release(); chirp.release(); bhat.release(); wpost.release(); N = 0; }
};
struct PoissonCache {                 // DCT matrices / FFT plans per image size
    int w = 0, h = 0;
    DevBuf Dw, DhT, work, work2, partial, dc;
    DctPlanBufs plan_w, plan_h;
    void release() { Dw.release(); DhT.release(); work.release(); work2.release(); partial.release(); dc.release(); plan_w.release(); plan_h.release(); w = h = 0; }
};

} // namespace gdpt

namespace gdpt {
// Queue buffers of one wave in flight.  Slot 0 is the context's own set (also used by the path integrator and the
// shift-mapped mode); slot 1 lets the next wave run on a second stream while the late, small bounces of the previous one
// leave most SMs idle.
struct WaveSlot {
    DevBuf *gen, *rays, *hits, *rec_a, *rec_b, *off, *counters, *sort_keys, *sort_order, *sort_hist, *res;
    cudaStream_t stream = nullptr;     // slot 0: the context's stream
    cudaEvent_t done = nullptr;        // the slot's last wave has been queued up to here
    cudaEvent_t reduced = nullptr;     // ... and its per-path results have been added to the images (the slot's buffers are free again)
    bool busy = false;                 // a wave of the current accumulation used the slot
};
// What the first (host-synchronised) wave of a configuration measured: live paths after the primary stage and after
// every bounce.  Later waves of the same configuration are launched from these numbers without reading anything back.
struct WaveCalibration {
    bool valid = false;
    long long n_paths = 0; int max_depth = 0, rr_depth = 0, precision = -1, width = 0, y0 = 0, y1 = 0, tail_paths = 0, strip_count = 0, strip_phase = 0;
    std::vector<int> counts;           // counts[0] after k_primary, counts[b] after bounce b
};
} // namespace gdpt

struct gdpt_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    int width = 0, height = 0;
    int max_depth = -1, rr_depth = 5, spp = 4;
    double bounds[4] = {0, 0, 0, 0};
    // scene arrays
    gdpt::DevBuf shapes, spheres, indices, materials, img3, img1;
    gdpt::DevBuf pos_d, nrm_d, uv_d, tex_d, texels3_d, texels1_d, light_d;   // exact mode
    gdpt::DevBuf pos_f, nrm_f, uv_f, tex_f, texels3_f, texels1_f, light_f;   // fast mode
    gdpt::DevBuf tri_recs, tri_recs_d;      // packed per-triangle vertex records (k_build_tri_recs), f32 and f64
    bool use_tri_recs = true;               // "tri_records" option
    gdpt::DevBuf lcdf_d, lcdf_f, lpmf_d, lpmf_f, tcdf_d, tcdf_f, area_d, area_f, lshape;   // light sampling tables (path integrator)
    bool has_envmap = false, envmap_ready = false;   // the scene has an environment map / its image and transforms were provided
    gdpt::DevBuf ecr_d, ecr_f, epr_d, epr_f, ecm_d, ecm_f, epm_d, epm_f;   // the environment map's TableDist2D
    gdpt::DevBuf shadow_rays, shadow_hits, prec_a, prec_b;                                  // path-integrator queues
    gdpt::DevBuf bvh_nodes, bvh_prims;
    gdpt::DScene<double> sc_d;
    gdpt::DScene<float> sc_f;
    gdpt::DBvh bvh;
    int bvh_depth = 0;
    // wavefront queues
    gdpt::DevBuf gen, rays, hits, rec_a, rec_b, off, counters, recout, xs, ys, streams;
    static constexpr int MAX_SLOTS = 4;
    gdpt::DevBuf res0;                      // per-path result records of slot 0 (deterministic accumulation)
    gdpt::DevBuf xbuf[MAX_SLOTS - 1][11];   // queue buffers of the wave slots beyond the first
    cudaStream_t reduce_stream = nullptr;   // k_wave_reduce of every wave, in wave order
    cudaEvent_t ev_reduce_join = nullptr;
    bool reduce_forked = false;             // reduce_stream has joined the context's stream for the current accumulation
    bool deterministic = true;              // per-path result records + ordered reduction instead of atomicAdd(double) ("deterministic")
    gdpt::WaveSlot slot[MAX_SLOTS];
    gdpt::WaveCalibration cal;
    bool async_waves = true;          // launch waves from the calibration without per-bounce host syncs (gdpt_set_option "async_waves")
    int wave_streams = 3;             // waves in flight on separate streams (1..MAX_SLOTS; "wave_streams"): 2 -> 3 is worth 1-2 %, 4 nothing more
    cudaEvent_t ev_fork = nullptr;
    unsigned forked = 0;              // bit k: slot k's stream has work of the current accumulation in flight
    int wave_index = 0;
    bool shade_sorted = false;        // k_bounce walks the queue in the traced (sorted) order instead of queue order ("shade_sorted")
    bool poisson_fft = true;          // the FFT form of the DCT-I (gdpt_poisson_fft.cuh); 0: the dense-matrix form ("poisson_fft")
    bool time_kernels = false;        // keep the host-synchronised wave loop with per-kernel event timing ("time_kernels"; bench.py's roofline pass)
    gdpt::DevBuf dev_stats;           // [0] rays traced by launches sized on the device (folded into stats.rays_traced by gdpt_get_stats)
    int *h_counts = nullptr;          // pinned
    // accumulation / output images
    gdpt::DevBuf img, cx0, cy0, cx1, cy1, cx, cy, out;
    gdpt::PoissonCache pc;
    gdpt_stats stats;
    cudaEvent_t ev[8];
    int sort_cell_bits = 5, sort_dir_bits = 3;   // ray-sort key: 2^5 cells per axis, 2^3 x 2^3 direction bins (GDPT_SORT_BITS="cell,dir"; gdpt_wavefront.cuh)
    bool sort_rays = true;            // re-order bounce-ray queues by origin cell + direction octant (GDPT_SORT_RAYS=0 disables)
    float box_lo[3] = {0, 0, 0}, box_hi[3] = {0, 0, 0};   // float scene box (BVH root)
    gdpt::DevBuf sort_keys, sort_order, sort_hist;
    int tail_paths = 1 << 16;         // finish a wave in one launch once this few paths are alive (GDPT_TAIL_PATHS)
    int shade_variant = 0;            // tuning knob (GDPT_SHADE_VARIANT): launch bounds of the diffuse-only float shade kernels
    int trace_variant = -1;           // tuning knob (GDPT_TRACE_VARIANT); -1 = the default kernel
    bool lambertian_only = false;     // every material is a Lambertian -> specialised shade kernels
    bool count_traversal = false;     // bench-only: count nodes/primitives per ray
    bool own_stream = true;
    bool bvh_pending = false;         // created with GDPT_CREATE_NO_BVH and not yet given a tree (gdpt_bvh_adopt)
    gdpt::DevBuf trav_counters;
};

namespace gdpt {

// Level 0 of every image is expanded, mip-mapped (make_mipmap, mipmap.h:28-44) and copied to float on the device:
// gdpt_textures.cu.
int build_texture_pools(gdpt_ctx *c, const gdpt_scene_desc *d, std::vector<DImage> &im3, std::vector<DImage> &im1);
// gdpt_bvh_gpu.cu: the tree built on the device (0: done; > 0: fall back to the host builder; < 0: error)
int build_bvh_device(gdpt_ctx *c, const std::vector<DShape> &shapes, double split_budget);

// level 0 of one image on the host (only the environment map's sampling table needs it there)
inline int decode_image(const gdpt_image &gi, int C, std::vector<double> &out) {
    if (gi.channels != C || gi.width <= 0 || gi.height <= 0) return fail(GDPT_ERR_ARG, "bad image (%d x %d x %d)", gi.width, gi.height, gi.channels);
    size_t n = (size_t)gi.width * gi.height * C;
    out.resize(n);
    if (gi.data) memcpy(out.data(), gi.data, n * sizeof(double));
    else if (gi.palette && gi.index) { for (size_t i = 0; i < n; i++) { if (gi.index[i] >= gi.palette_size) return fail(GDPT_ERR_ARG, "palette index out of range"); out[i] = gi.palette[gi.index[i]]; } }
    else return fail(GDPT_ERR_ARG, "image has neither data nor palette");
    return 0;
}

template <typename R> void fill_scene_common(DScene<R> &s, const gdpt_scene_desc *d) {
    for (int i = 0; i < 16; i++) { s.s2c[i] = (R)d->sample_to_cam[i]; s.c2w[i] = (R)d->cam_to_world[i]; }
    s.width = d->width; s.height = d->height; s.filter_kind = d->filter_kind; s.filter_param = (R)d->filter_param;
    s.n_shapes = d->n_shapes; s.n_materials = d->n_materials; s.n_lights = d->n_lights;
    s.max_depth = d->max_depth; s.rr_depth = d->rr_depth;
}

inline int create_context(const gdpt_scene_desc *d, int device, int flags, gdpt_ctx **out) {
    if (!d || !out) return fail(GDPT_ERR_ARG, "null argument");
    if (d->width <= 1 || d->height <= 1) return fail(GDPT_ERR_ARG, "film must be at least 2x2");
    if (d->n_positions < 0 || d->n_triangles < 0 || d->n_normals < 0 || d->n_uvs < 0 || d->n_positions > 0x7fffffffLL / 4 || d->n_triangles > 0x7fffffffLL / 4 ||
        d->n_normals > 0x7fffffffLL / 4 || d->n_uvs > 0x7fffffffLL / 4 || d->n_shapes < 0 || d->n_materials < 0 || d->n_textures < 0 || d->n_lights < 0)
        return fail(GDPT_ERR_ARG, "pool sizes out of range");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(GDPT_ERR_CUDA, "no CUDA device: libgdpt has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(GDPT_ERR_ARG, "device %d out of range (%d devices)", device, ndev);
    GD_CUDA(cudaSetDevice(device));
    setup_pool(device);
    gdpt_ctx *c = new gdpt_ctx();
    c->device = device; c->width = d->width; c->height = d->height;
    c->max_depth = d->max_depth; c->rr_depth = d->rr_depth; c->spp = d->samples_per_pixel;
    memset(&c->stats, 0, sizeof(c->stats));
    *out = c;
    GD_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    for (auto &e : c->ev) GD_CUDA(cudaEventCreate(&e));
    GD_CUDA(cudaMallocHost((void **)&c->h_counts, 64));
    GD_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    c->slot[0] = gdpt::WaveSlot{&c->gen, &c->rays, &c->hits, &c->rec_a, &c->rec_b, &c->off, &c->counters, &c->sort_keys, &c->sort_order, &c->sort_hist, &c->res0};
    GD_CUDA(cudaStreamCreateWithFlags(&c->reduce_stream, cudaStreamNonBlocking));
    GD_CUDA(cudaEventCreateWithFlags(&c->ev_reduce_join, cudaEventDisableTiming));
    if (const char *e = getenv("GDPT_DETERMINISTIC")) c->deterministic = atoi(e) != 0;
    for (int k = 1; k < gdpt_ctx::MAX_SLOTS; k++) {
        gdpt::DevBuf *b = c->xbuf[k - 1];
        c->slot[k] = gdpt::WaveSlot{b, b + 1, b + 2, b + 3, b + 4, b + 5, b + 6, b + 7, b + 8, b + 9, b + 10};
        GD_CUDA(cudaStreamCreateWithFlags(&c->slot[k].stream, cudaStreamNonBlocking));
    }
    for (auto &sl : c->slot) { GD_CUDA(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming)); GD_CUDA(cudaEventCreateWithFlags(&sl.reduced, cudaEventDisableTiming)); }
    if (int r = c->dev_stats.ensure(64)) return r;
    GD_CUDA(cudaMemset(c->dev_stats.p, 0, 64));
    if (const char *e = getenv("GDPT_ASYNC_WAVES")) c->async_waves = atoi(e) != 0;
    if (const char *e = getenv("GDPT_POISSON_FFT")) c->poisson_fft = atoi(e) != 0;
    if (const char *e = getenv("GDPT_WAVE_STREAMS")) c->wave_streams = std::min(gdpt_ctx::MAX_SLOTS, std::max(1, atoi(e)));

    const bool verbose = getenv("GDPT_VERBOSE") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto lap = [&](const char *what) {
        auto now = std::chrono::steady_clock::now();
        if (verbose) fprintf(stderr, "[gdpt_create] %-18s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
    lap("cuda init");
    // ---- shapes + BVH
    std::vector<DShape> shapes(d->n_shapes);
    std::vector<HostShape> hs(d->n_shapes);
    for (int i = 0; i < d->n_shapes; i++) {
        const int32_t *s = d->shapes + 6 * i; const int64_t *o = d->shape_offsets + 4 * i;
        DShape &ds = shapes[i];
        ds.kind = s[0]; ds.material_id = s[1]; ds.area_light_id = s[2]; ds.flags = s[5];
        ds.vtx_off = (int)o[0]; ds.tri_off = (int)o[1]; ds.nrm_off = (int)o[2]; ds.uv_off = (int)o[3];
        ds.n_tris = s[4]; ds.cdf_off = -1;
        // every range the builder and the shading kernels will index must lie inside the pools
        if (s[0] != 0 && s[0] != 1) return fail(GDPT_ERR_ARG, "shape %d: unknown kind %d", i, s[0]);
        for (int k = 0; k < 4; k++) if (o[k] < 0 || o[k] > 0x7fffffffLL / 4) return fail(GDPT_ERR_ARG, "shape %d: pool offset %lld out of range", i, (long long)o[k]);
        if (s[0] == 1) {
            if (s[3] < 0 || s[4] < 0 || o[0] + s[3] > d->n_positions || o[1] + s[4] > d->n_triangles) return fail(GDPT_ERR_ARG, "shape %d: vertex / triangle range outside the pools", i);
            if ((s[5] & 1) && o[2] + s[3] > d->n_normals) return fail(GDPT_ERR_ARG, "shape %d: normal range outside the pool", i);
            if ((s[5] & 2) && o[3] + s[3] > d->n_uvs) return fail(GDPT_ERR_ARG, "shape %d: uv range outside the pool", i);
        }
        if (ds.material_id < 0 || ds.material_id >= d->n_materials) return fail(GDPT_ERR_ARG, "shape %d: material id %d out of range", i, ds.material_id);
        if (ds.area_light_id >= d->n_lights) return fail(GDPT_ERR_ARG, "shape %d: light id out of range", i);
        HostShape &h = hs[i];
        h.kind = s[0]; h.nverts = s[3]; h.ntris = s[4];
        h.positions = d->positions + 3 * o[0]; h.indices = d->indices + 3 * o[1];
        for (int k = 0; k < 4; k++) h.sphere[k] = d->spheres[4 * i + k];
        if (h.kind == 1) for (int t = 0; t < 3 * h.ntris; t++) if (h.indices[t] < 0 || h.indices[t] >= h.nverts) return fail(GDPT_ERR_ARG, "shape %d: vertex index out of range", i);
    }
    // the BVH is built on its own host thread(s) while this one decodes textures and uploads the pools
    Bvh bvh;
    const bool no_bvh = (flags & GDPT_CREATE_NO_BVH) != 0;     // the tree comes from another process: only the bounds are needed here
    // Two builders (Embree's build-quality levels, scene.cpp:22, in this library's terms): the host's binned SAH with triangle
    // pre-splitting (gdpt_bvh.cpp: ~45 ms for Sponza, the better tree: 17.4 nodes + 2.8 primitives per ray) and the device's PLOC
    // (gdpt_bvh_gpu.cu: ~9 ms, 18.1 nodes + 4.7 primitives per ray, 5-6 % slower rendering).  Without a flag the planned work
    // decides: the host build pays for itself from about 1e8 pixel-samples (Sponza 1080p at ~50 spp on one GPU:
    // 5.5 % of the render time against ~35 ms of build time).
    const char *bvh_mode = getenv("GDPT_BVH");
    const double planned = (double)d->width * d->height * std::max(1, d->samples_per_pixel);
    bool device_bvh = (flags & GDPT_CREATE_BVH_DEVICE) ? true : (flags & GDPT_CREATE_BVH_HOST) ? false : planned < 1e8;
    if (bvh_mode) device_bvh = std::string(bvh_mode) != "host";
    if (no_bvh) device_bvh = false;
    std::thread bvh_thread([&] { if (no_bvh || device_bvh) scene_bounds(hs, bvh.lo, bvh.hi); else build_bvh(hs, bvh); });
    struct Joiner { std::thread &t; ~Joiner() { if (t.joinable()) t.join(); } } bvh_joiner{bvh_thread};
    // ---- geometry pools
    std::vector<double> pos(d->positions, d->positions + 3 * (size_t)d->n_positions), nrm(d->normals, d->normals + 3 * (size_t)d->n_normals),
        uv(d->uvs, d->uvs + 2 * (size_t)d->n_uvs);
    std::vector<int> idx(d->indices, d->indices + 3 * (size_t)d->n_triangles);
    auto to_f = [](const std::vector<double> &v) { return std::vector<float>(v.begin(), v.end()); };
    if (int r = upload(c->pos_d, pos)) return r; if (int r = upload(c->nrm_d, nrm)) return r; if (int r = upload(c->uv_d, uv)) return r;
    if (int r = upload(c->pos_f, to_f(pos))) return r; if (int r = upload(c->nrm_f, to_f(nrm))) return r; if (int r = upload(c->uv_f, to_f(uv))) return r;
    if (int r = upload(c->indices, idx)) return r;

    lap("geometry upload");
    if (device_bvh) {
        if (int r = upload(c->shapes, shapes)) return r;          // (uploaded again below once the light tables have set cdf_off)
        { std::vector<double> sp(d->spheres, d->spheres + 4 * (size_t)d->n_shapes); if (int r = upload(c->spheres, sp)) return r; }
        c->bvh.spheres = c->spheres.as<double>();
        double split_budget = 0.0;      // the device pre-split exists (k_bvh_split) but does not pay under bottom-up clustering: DESIGN.md 4.1
        if (const char *e = getenv("GDPT_BVH_SPLIT_DEVICE")) split_budget = atof(e);
        const int rb = build_bvh_device(c, shapes, split_budget);
        if (rb < 0) return rb;
        if (rb > 0) { device_bvh = false; bvh_thread.join(); build_bvh(hs, bvh); }     // too deep for the traversal stack: host builder
        lap("bvh build (device)");
    }
    // ---- materials / textures / images
    std::vector<DMaterial> mats(d->n_materials);
    for (int i = 0; i < d->n_materials; i++) {
        const double *m = d->materials + 14 * i;
        mats[i].type = (int)m[0]; mats[i].eta = m[1];
        if (mats[i].type < 0 || mats[i].type >= MAT_COUNT) return fail(GDPT_ERR_ARG, "material %d: unknown type %d", i, mats[i].type);
        for (int k = 0; k < 12; k++) { mats[i].tex[k] = (int)m[2 + k]; if (mats[i].tex[k] >= d->n_textures) return fail(GDPT_ERR_ARG, "material %d: texture id out of range", i); }
    }
    if (const char *tv = getenv("GDPT_TRACE_VARIANT")) c->trace_variant = atoi(tv);
    if (const char *sv = getenv("GDPT_SORT_RAYS")) c->sort_rays = atoi(sv) != 0;
    if (const char *tp = getenv("GDPT_TAIL_PATHS")) c->tail_paths = atoi(tp);
    if (const char *sb = getenv("GDPT_SORT_BITS")) {
        int cb = 0, db = 0;
        if (sscanf(sb, "%d,%d", &cb, &db) == 2 && cb >= 2 && db >= 0 && 3 * cb + 2 * db >= 15 && 3 * cb + 2 * db <= 25) { c->sort_cell_bits = cb; c->sort_dir_bits = db; }
    }
    if (const char *sv = getenv("GDPT_SHADE_VARIANT")) c->shade_variant = atoi(sv);
    c->lambertian_only = !mats.empty();
    for (const DMaterial &m : mats) if (m.type != MAT_LAMBERTIAN) c->lambertian_only = false;
    // Scenes with more than diffuse surfaces shade in the TRACED order (the ray sort's permutation: origin cell + direction
    // bin): neighbouring hits share meshes, textures and mostly the material, so a warp runs one arm of the nine-way BSDF
    // dispatch instead of several — the north star's material-sorted queue, obtained from the spatial sort that exists anyway.
    // B200: Disney config 528 -> 881 Mpaths/s; all-Lambertian Sponza loses 8 % to the gathered record loads, so it keeps queue order.
    c->shade_sorted = !c->lambertian_only;
    if (const char *e = getenv("GDPT_SHADE_SORTED")) c->shade_sorted = atoi(e) != 0;
    if (int r = upload(c->materials, mats)) return r;
    std::vector<DTexture<double>> td(d->n_textures); std::vector<DTexture<float>> tf(d->n_textures);
    for (int i = 0; i < d->n_textures; i++) {
        const double *t = d->textures + 14 * i;
        td[i].kind = (int)t[0]; td[i].image_id = (int)t[8];
        bool spec = t[1] != 0;
        if (td[i].kind == TEX_IMAGE && (td[i].image_id < 0 || td[i].image_id >= (spec ? d->n_images3 : d->n_images1))) return fail(GDPT_ERR_ARG, "texture %d: image id out of range", i);
        for (int k = 0; k < 3; k++) { td[i].v0[k] = t[2 + k]; td[i].v1[k] = t[5 + k]; }
        td[i].uscale = t[9]; td[i].vscale = t[10]; td[i].uoffset = t[11]; td[i].voffset = t[12];
        tf[i].kind = td[i].kind; tf[i].image_id = td[i].image_id;
        for (int k = 0; k < 3; k++) { tf[i].v0[k] = (float)td[i].v0[k]; tf[i].v1[k] = (float)td[i].v1[k]; }
        tf[i].uscale = (float)t[9]; tf[i].vscale = (float)t[10]; tf[i].uoffset = (float)t[11]; tf[i].voffset = (float)t[12];
    }
    if (int r = upload(c->tex_d, td)) return r; if (int r = upload(c->tex_f, tf)) return r;
    std::vector<DImage> im3, im1;
    if (int r = build_texture_pools(c, d, im3, im1)) return r;
    lap("textures (device)");
    if (int r = upload(c->img3, im3)) return r; if (int r = upload(c->img1, im1)) return r;
    std::vector<double> li(3 * (size_t)std::max(d->n_lights, 1), 0.0);
    for (int i = 0; i < d->n_lights; i++) for (int k = 0; k < 3; k++) li[3 * i + k] = d->lights[5 * i + 2 + k];
    if (int r = upload(c->light_d, li)) return r; if (int r = upload(c->light_f, to_f(li))) return r;
    // light sampling tables, all as make_table_dist_1d builds them (table_dist.cpp:3-26)
    auto table = [](const std::vector<double> &f, std::vector<double> &pmf, std::vector<double> &cdf) {
        pmf = f; cdf.assign(f.size() + 1, 0.0);
        for (size_t i = 0; i < f.size(); i++) cdf[i + 1] = cdf[i] + pmf[i];
        double total = cdf.back();
        if (total > 0) { for (size_t i = 0; i < pmf.size(); i++) { pmf[i] /= total; cdf[i] /= total; } }   // the last cdf entry keeps the total
        else { for (size_t i = 0; i < pmf.size(); i++) { pmf[i] = 1.0 / (double)pmf.size(); cdf[i] = (double)i / (double)pmf.size(); } if (!cdf.empty()) cdf.back() = 1; }
    };
    std::vector<double> shape_area;
    {   // init_sampling_dist of the emissive meshes (triangle_mesh.inl:60-75) and the surface areas
        std::vector<double> tri_cdf, area(std::max(d->n_shapes, 1), 0.0);
        for (int i = 0; i < d->n_shapes; i++) {
            const HostShape &h = hs[i];
            if (h.kind == 0) { area[i] = 4 * 3.14159265358979323846 * h.sphere[3] * h.sphere[3]; continue; }
            if (shapes[i].area_light_id < 0) continue;
            std::vector<double> ta(h.ntris), pm, cd; double total = 0;
            for (int t = 0; t < h.ntris; t++) {
                const double *v0 = h.positions + 3 * h.indices[3 * t], *v1 = h.positions + 3 * h.indices[3 * t + 1], *v2 = h.positions + 3 * h.indices[3 * t + 2];
                double e1[3] = {v1[0] - v0[0], v1[1] - v0[1], v1[2] - v0[2]}, e2[3] = {v2[0] - v0[0], v2[1] - v0[1], v2[2] - v0[2]};
                double cx = e1[1] * e2[2] - e1[2] * e2[1], cy = e1[2] * e2[0] - e1[0] * e2[2], cz = e1[0] * e2[1] - e1[1] * e2[0];
                ta[t] = sqrt(cx * cx + cy * cy + cz * cz) / 2; total += ta[t];
            }
            table(ta, pm, cd);
            shapes[i].cdf_off = (int)tri_cdf.size(); area[i] = total;
            tri_cdf.insert(tri_cdf.end(), cd.begin(), cd.end());
        }
        if (tri_cdf.empty()) tri_cdf.push_back(0.0);
        if (int r = upload(c->tcdf_d, tri_cdf)) return r; if (int r = upload(c->tcdf_f, to_f(tri_cdf))) return r;
        if (int r = upload(c->area_d, area)) return r; if (int r = upload(c->area_f, to_f(area))) return r;
        shape_area = area;
    }

    lap("light tables");
    if (bvh_thread.joinable()) bvh_thread.join();
    lap("bvh build (rest)");
    if (!device_bvh) {
        if (3 * bvh.max_depth + 1 > STACK_TOTAL) return fail(GDPT_ERR_UNSUPPORTED, "BVH depth %d exceeds the traversal stack", bvh.max_depth);
        c->bvh_depth = bvh.max_depth;
    }
    for (int k = 0; k < 3; k++) { c->box_lo[k] = bvh.lo[k]; c->box_hi[k] = bvh.hi[k]; }
    if (int r = upload(c->shapes, shapes)) return r;
    { std::vector<double> sp(d->spheres, d->spheres + 4 * (size_t)d->n_shapes); if (int r = upload(c->spheres, sp)) return r; }
    c->bvh.spheres = c->spheres.as<double>();
    if (!device_bvh) {
        if (int r = upload(c->bvh_nodes, bvh.nodes)) return r;
        if (int r = upload(c->bvh_prims, bvh.prims)) return r;
        c->bvh.nodes = c->bvh_nodes.as<float4>(); c->bvh.prims = c->bvh_prims.as<float4>();
        c->bvh.n_nodes = bvh.prims.empty() ? 0 : (int)bvh.nodes.size(); c->bvh.n_prims = (int)bvh.prims.size();
    }
    c->bvh_pending = no_bvh;
    // bounds sphere (scene.cpp:30-34): float box from the "Embree" scene, arithmetic in double
    {
        double lb[3] = {bvh.lo[0], bvh.lo[1], bvh.lo[2]}, ub[3] = {bvh.hi[0], bvh.hi[1], bvh.hi[2]};
        double dx = ub[0] - lb[0], dy = ub[1] - lb[1], dz = ub[2] - lb[2];
        c->bounds[0] = sqrt(dx * dx + dy * dy + dz * dz) / 2;
        for (int k = 0; k < 3; k++) c->bounds[1 + k] = (lb[k] + ub[k]) / 2.0;
    }
    double eps = std::min(c->bounds[0] * 1e-5, 0.01);   // scene.h:104-106
    fill_scene_common(c->sc_d, d); fill_scene_common(c->sc_f, d);
    {   // light_power (diffuse_area_light.inl:1-3, envmap.inl:1-5) and the light distribution (scene.cpp:46-52); the
        // environment map's TableDist2D over luminance * sin(elevation) of its level-0 texels (envmap.inl:76-98)
        std::vector<double> power(d->n_lights), lpmf, lcdf; std::vector<int> lshape(std::max(d->n_lights, 1), -1);
        std::vector<double> cdf_rows(1, 0.0), pdf_rows(1, 0.0), cdf_marg(1, 0.0), pdf_marg(1, 0.0);
        c->has_envmap = false; c->envmap_ready = false;
        c->sc_d.env_light = c->sc_f.env_light = -1; c->sc_d.env_tex = c->sc_f.env_tex = -1;
        c->sc_d.env_w = c->sc_f.env_w = 0; c->sc_d.env_h = c->sc_f.env_h = 0;
        for (int i = 0; i < d->n_lights; i++) {
            const double *l = d->lights + 5 * i;
            if ((int)l[0] != 0) {
                c->has_envmap = true; power[i] = 0;
                const int tid = (int)l[1];
                if (tid < 0 || !d->envmap_to_world || !d->envmap_to_local) continue;          // not provided: path refuses the scene
                if (tid >= d->n_textures || td[tid].kind != TEX_IMAGE || d->textures[14 * tid + 1] == 0) return fail(GDPT_ERR_UNSUPPORTED, "environment map %d is not an RGB image texture", i);
                const DImage &im = im3[td[tid].image_id];
                const int w = im.w[0], h = im.h[0];
                std::vector<double> env0;
                if (int r = decode_image(d->images3[td[tid].image_id], 3, env0)) return r;
                const double *tex = env0.data();
                auto imod = [](int a, int b) { int r = a % b; return r < 0 ? r + b : r; };
                std::vector<double> f((size_t)w * h);
                for (int y = 0; y < h; y++) {
                    double v = (y + 0.5) / (double)h, sin_el = sin(3.14159265358979323846 * v);
                    for (int x = 0; x < w; x++) {
                        double u = (x + 0.5) / (double)w;
                        // lookup(mipmap, u, v, 0): bilinear at level 0 (mipmap.h:46-69)
                        double uu = u * w - 0.5, vv = v * h - 0.5;
                        int ufi = imod((int)uu, w), vfi = imod((int)vv, h), uci = imod(ufi + 1, w), vci = imod(vfi + 1, h);
                        double uo = uu - ufi, vo = vv - vfi, lum[4];
                        const int xs[4] = {ufi, ufi, uci, uci}, ys[4] = {vfi, vci, vfi, vci};
                        double rgb[3];
                        for (int ch = 0; ch < 3; ch++) {
                            double ff = tex[3 * ((size_t)ys[0] * w + xs[0]) + ch], fc = tex[3 * ((size_t)ys[1] * w + xs[1]) + ch];
                            double cf = tex[3 * ((size_t)ys[2] * w + xs[2]) + ch], cc = tex[3 * ((size_t)ys[3] * w + xs[3]) + ch];
                            rgb[ch] = ff * (1 - uo) * (1 - vo) + fc * (1 - uo) * vo + cf * uo * (1 - vo) + cc * uo * vo;
                        }
                        (void)lum;
                        f[(size_t)y * w + x] = (rgb[0] * 0.212671 + rgb[1] * 0.715160 + rgb[2] * 0.072169) * sin_el;
                    }
                }
                // make_table_dist_2d (table_dist.cpp:45-123)
                cdf_rows.assign((size_t)h * (w + 1), 0.0); pdf_rows.assign((size_t)h * w, 0.0);
                for (int y = 0; y < h; y++) {
                    double *cr = cdf_rows.data() + (size_t)y * (w + 1), *pr = pdf_rows.data() + (size_t)y * w;
                    for (int x = 0; x < w; x++) cr[x + 1] = cr[x] + f[(size_t)y * w + x];
                    double integral = cr[w];
                    if (integral > 0) { for (int x = 0; x < w; x++) cr[x] /= integral; for (int x = 0; x < w; x++) pr[x] = f[(size_t)y * w + x] / integral; }
                    else { for (int x = 0; x < w; x++) { pr[x] = 1.0 / (double)w; cr[x] = (double)x / (double)w; } cr[w] = 1; }
                }
                cdf_marg.assign((size_t)h + 1, 0.0); pdf_marg.assign((size_t)h, 0.0);
                for (int y = 0; y < h; y++) cdf_marg[y + 1] = cdf_marg[y] + cdf_rows[(size_t)y * (w + 1) + w];
                const double total = cdf_marg.back();
                if (total > 0) {
                    for (int y = 0; y < h; y++) cdf_marg[y] /= total;
                    cdf_marg[h] = 1;
                    for (int y = 0; y < h; y++) pdf_marg[y] = cdf_rows[(size_t)y * (w + 1) + w] / total;
                } else {
                    for (int y = 0; y < h; y++) { pdf_marg[y] = 1.0 / (double)h; cdf_marg[y] = (double)y / (double)h; }
                    cdf_marg[h] = 1;
                }
                for (int y = 0; y < h; y++) cdf_rows[(size_t)y * (w + 1) + w] = 1;
                power[i] = 3.14159265358979323846 * c->bounds[0] * c->bounds[0] * total / (double)(w * h);
                c->envmap_ready = true;
                c->sc_d.env_light = c->sc_f.env_light = i; c->sc_d.env_tex = c->sc_f.env_tex = tid;
                c->sc_d.env_w = c->sc_f.env_w = w; c->sc_d.env_h = c->sc_f.env_h = h;
                c->sc_d.env_scale = l[2]; c->sc_f.env_scale = (float)l[2];
                for (int k = 0; k < 16; k++) {
                    c->sc_d.env_to_world[k] = d->envmap_to_world[k]; c->sc_f.env_to_world[k] = (float)d->envmap_to_world[k];
                    c->sc_d.env_to_local[k] = d->envmap_to_local[k]; c->sc_f.env_to_local[k] = (float)d->envmap_to_local[k];
                }
                continue;
            }
            int sid = (int)l[1];
            if (sid < 0 || sid >= d->n_shapes) return fail(GDPT_ERR_ARG, "light %d: shape id out of range", i);
            lshape[i] = sid;
            power[i] = (l[2] * 0.212671 + l[3] * 0.715160 + l[4] * 0.072169) * shape_area[sid] * 3.14159265358979323846;
        }
        table(power, lpmf, lcdf);
        if (lpmf.empty()) lpmf.push_back(0.0);
        if (int r = upload(c->lcdf_d, lcdf)) return r; if (int r = upload(c->lcdf_f, to_f(lcdf))) return r;
        if (int r = upload(c->lpmf_d, lpmf)) return r; if (int r = upload(c->lpmf_f, to_f(lpmf))) return r;
        if (int r = upload(c->lshape, lshape)) return r;
        if (int r = upload(c->ecr_d, cdf_rows)) return r; if (int r = upload(c->ecr_f, to_f(cdf_rows))) return r;
        if (int r = upload(c->epr_d, pdf_rows)) return r; if (int r = upload(c->epr_f, to_f(pdf_rows))) return r;
        if (int r = upload(c->ecm_d, cdf_marg)) return r; if (int r = upload(c->ecm_f, to_f(cdf_marg))) return r;
        if (int r = upload(c->epm_d, pdf_marg)) return r; if (int r = upload(c->epm_f, to_f(pdf_marg))) return r;
        c->sc_d.env_cdf_rows = c->ecr_d.as<double>(); c->sc_f.env_cdf_rows = c->ecr_f.as<float>();
        c->sc_d.env_pdf_rows = c->epr_d.as<double>(); c->sc_f.env_pdf_rows = c->epr_f.as<float>();
        c->sc_d.env_cdf_marg = c->ecm_d.as<double>(); c->sc_f.env_cdf_marg = c->ecm_f.as<float>();
        c->sc_d.env_pdf_marg = c->epm_d.as<double>(); c->sc_f.env_pdf_marg = c->epm_f.as<float>();
    }

    lap("bvh upload");
    c->sc_d.eps = eps; c->sc_f.eps = (float)eps;
    c->sc_d.shapes = c->sc_f.shapes = c->shapes.as<DShape>();
    c->sc_d.spheres = c->sc_f.spheres = c->spheres.as<double>();
    c->sc_d.indices = c->sc_f.indices = c->indices.as<int>();
    c->sc_d.materials = c->sc_f.materials = c->materials.as<DMaterial>();
    c->sc_d.img3 = c->sc_f.img3 = c->img3.as<DImage>(); c->sc_d.img1 = c->sc_f.img1 = c->img1.as<DImage>();
    c->sc_d.positions = c->pos_d.as<double>(); c->sc_d.normals = c->nrm_d.as<double>(); c->sc_d.uvs = c->uv_d.as<double>();
    c->sc_f.positions = c->pos_f.as<float>(); c->sc_f.normals = c->nrm_f.as<float>(); c->sc_f.uvs = c->uv_f.as<float>();
    c->sc_d.tri_recs = nullptr; c->sc_f.tri_recs = nullptr;
    if (d->n_triangles > 0) {           // a hit triangle's vertex data is read from one packed record (96 B in f32, 192 B in f64)
        if (int r = c->tri_recs.ensure((size_t)d->n_triangles * 24 * sizeof(float))) return r;
        if (int r = c->tri_recs_d.ensure((size_t)d->n_triangles * 24 * sizeof(double))) return r;
        const unsigned nb = (unsigned)((d->n_triangles + 255) / 256);
        k_build_tri_recs<float><<<nb, 256, 0, c->stream>>>(d->n_shapes, c->shapes.as<DShape>(), c->indices.as<int>(), c->pos_f.as<float>(), c->nrm_f.as<float>(),
                                                          c->uv_f.as<float>(), (long long)d->n_triangles, c->tri_recs.as<float>());
        k_build_tri_recs<double><<<nb, 256, 0, c->stream>>>(d->n_shapes, c->shapes.as<DShape>(), c->indices.as<int>(), c->pos_d.as<double>(), c->nrm_d.as<double>(),
                                                           c->uv_d.as<double>(), (long long)d->n_triangles, c->tri_recs_d.as<double>());
        GD_CUDA(cudaGetLastError());
        c->stats.kernel_launches += 2;
        if (c->use_tri_recs) { c->sc_f.tri_recs = c->tri_recs.as<float>(); c->sc_d.tri_recs = c->tri_recs_d.as<double>(); }
    }
    c->sc_d.textures = c->tex_d.as<DTexture<double>>(); c->sc_f.textures = c->tex_f.as<DTexture<float>>();
    c->sc_d.texels3 = c->texels3_d.as<double>(); c->sc_d.texels1 = c->texels1_d.as<double>();
    c->sc_f.texels3 = c->texels3_f.as<float4>(); c->sc_f.texels1 = c->texels1_f.as<float>();
    c->sc_d.light_intensity = c->light_d.as<double>(); c->sc_f.light_intensity = c->light_f.as<float>();
    c->sc_d.light_cdf = c->lcdf_d.as<double>(); c->sc_f.light_cdf = c->lcdf_f.as<float>();
    c->sc_d.light_pmf = c->lpmf_d.as<double>(); c->sc_f.light_pmf = c->lpmf_f.as<float>();
    c->sc_d.tri_cdf = c->tcdf_d.as<double>(); c->sc_f.tri_cdf = c->tcdf_f.as<float>();
    c->sc_d.shape_area = c->area_d.as<double>(); c->sc_f.shape_area = c->area_f.as<float>();
    c->sc_d.light_shape = c->sc_f.light_shape = c->lshape.as<int>();
    GD_CUDA(cudaDeviceSynchronize());   // pageable uploads above went through the legacy stream; the context's stream does not wait on it
    return 0;
}

// after the last wave of an accumulation: the context's stream waits for the other wave streams
inline int join_wave_streams(gdpt_ctx *c) {
    for (int k = 1; k < gdpt_ctx::MAX_SLOTS; k++) if (c->forked >> k & 1) {
        GD_CUDA(cudaEventRecord(c->slot[k].done, c->slot[k].stream));
        GD_CUDA(cudaStreamWaitEvent(c->stream, c->slot[k].done, 0));
    }
    c->forked = 0;
    if (c->reduce_forked) {
        GD_CUDA(cudaEventRecord(c->ev_reduce_join, c->reduce_stream));
        GD_CUDA(cudaStreamWaitEvent(c->stream, c->ev_reduce_join, 0));
        c->reduce_forked = false;
    }
    for (auto &sl : c->slot) sl.busy = false;
    c->wave_index = 0;
    return 0;
}

inline void destroy_context(gdpt_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    gdpt::DevBuf *all[] = {&c->shapes, &c->spheres, &c->indices, &c->materials, &c->img3, &c->img1, &c->pos_d, &c->nrm_d, &c->uv_d, &c->tex_d,
                           &c->texels3_d, &c->texels1_d, &c->light_d, &c->pos_f, &c->nrm_f, &c->uv_f, &c->tri_recs, &c->tri_recs_d, &c->tex_f, &c->texels3_f, &c->texels1_f,
                           &c->light_f, &c->lcdf_d, &c->lcdf_f, &c->lpmf_d, &c->lpmf_f, &c->tcdf_d, &c->tcdf_f, &c->area_d, &c->area_f, &c->lshape, &c->ecr_d, &c->ecr_f, &c->epr_d, &c->epr_f, &c->ecm_d, &c->ecm_f, &c->epm_d, &c->epm_f, &c->shadow_rays, &c->shadow_hits, &c->prec_a, &c->prec_b, &c->bvh_nodes, &c->bvh_prims, &c->gen, &c->rays, &c->hits, &c->rec_a, &c->rec_b, &c->off, &c->counters,
                           &c->recout, &c->xs, &c->ys, &c->streams, &c->trav_counters, &c->sort_keys, &c->sort_order, &c->sort_hist, &c->dev_stats, &c->res0, &c->img, &c->cx0, &c->cy0, &c->cx1, &c->cy1, &c->cx, &c->cy, &c->out};
    for (auto *b : all) b->release();
    for (auto &row : c->xbuf) for (auto &b : row) b.release();
    c->pc.release();
    if (c->h_counts) cudaFreeHost(c->h_counts);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    for (auto &sl : c->slot) { if (sl.done) cudaEventDestroy(sl.done); if (sl.reduced) cudaEventDestroy(sl.reduced); }
    if (c->ev_reduce_join) cudaEventDestroy(c->ev_reduce_join);
    if (c->reduce_stream) cudaStreamDestroy(c->reduce_stream);
    for (int k = 1; k < gdpt_ctx::MAX_SLOTS; k++) if (c->slot[k].stream) cudaStreamDestroy(c->slot[k].stream);
    for (auto &e : c->ev) if (e) cudaEventDestroy(e);
    if (c->stream && c->own_stream) cudaStreamDestroy(c->stream);
    delete c;
}

} // namespace gdpt
