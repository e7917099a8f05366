// C ABI of libgdpt.so (include/gdpt.h).  Host glue only: every arithmetic step of the path
// runs in the CUDA kernels of this directory (wavefront: gdpt_wave_f64.cu / _f32.cu, Poisson:
// gdpt_poisson.cu, validation launchers: gdpt_validate_*.cu).
#include "gdpt_context.cuh"
#include "gdpt_wavefront.cuh"

namespace gdpt {
// other translation units
template <typename R> int run_wave(gdpt_ctx *c, DScene<R> sc, WaveParams wp);
extern template int run_wave<double>(gdpt_ctx *, DScene<double>, WaveParams);
extern template int run_wave<float>(gdpt_ctx *, DScene<float>, WaveParams);
template <typename R> int run_shift_wave(gdpt_ctx *c, DScene<R> sc, WaveParams wp);
extern template int run_shift_wave<double>(gdpt_ctx *, DScene<double>, WaveParams);
extern template int run_shift_wave<float>(gdpt_ctx *, DScene<float>, WaveParams);
template <typename R> int run_path_wave(gdpt_ctx *c, DScene<R> sc, WaveParams wp);
extern template int run_path_wave<double>(gdpt_ctx *, DScene<double>, WaveParams);
extern template int run_path_wave<float>(gdpt_ctx *, DScene<float>, WaveParams);
template <typename R> int dbg_sample_light(gdpt_ctx *ctx, const DScene<R> &sc, int n, const double *ref_points, const double *rnd, double *out);
extern template int dbg_sample_light<double>(gdpt_ctx *, const DScene<double> &, int, const double *, const double *, double *);
extern template int dbg_sample_light<float>(gdpt_ctx *, const DScene<float> &, int, const double *, const double *, double *);
int dbg_occluded(gdpt_ctx *ctx, bool exact, int n, const double *rays, int32_t *out);
template <typename R> int dbg_aux(gdpt_ctx *ctx, const DScene<R> &sc, int kind, double *out);
extern template int dbg_aux<double>(gdpt_ctx *, const DScene<double> &, int, double *);
extern template int dbg_aux<float>(gdpt_ctx *, const DScene<float> &, int, double *);
int poisson_device(gdpt_ctx *c, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, double alpha, double *d_out);
int merge_device(gdpt_ctx *ctx, int w, int h, int ya, int yb, const double *d_cx0, const double *d_cx1, const double *d_cy0, const double *d_cy1, double *d_cx, double *d_cy);
int poisson_dist_rows_forward(gdpt_ctx *c, const gdpt_dist_geom *g, const double *d_u, const double *d_gx, const double *d_gy, double alpha,
                              double *d_send, double *d_dc_partial);
int poisson_dist_columns(gdpt_ctx *c, const gdpt_dist_geom *g, double alpha, const double *d_dc, const double *d_recv, double *d_send_back);
int poisson_dist_rows_inverse(gdpt_ctx *c, const gdpt_dist_geom *g, const double *d_recv_back, double *d_out);
template <typename R> int dbg_sample_primary(gdpt_ctx *ctx, const DScene<R> &sc, int n, const double *screen_pos, double *rays);
template <typename R> int dbg_intersect(gdpt_ctx *ctx, const DScene<R> &sc, int n, const double *rays, const double *rdiff, int32_t *ids, double *vtx);
template <typename R> int dbg_bsdf(gdpt_ctx *ctx, const DScene<R> &sc, int mat_id, int n, const double *vtx, const double *dirs, const double *rnd, double *out);
extern template int dbg_sample_primary<double>(gdpt_ctx *, const DScene<double> &, int, const double *, double *);
extern template int dbg_sample_primary<float>(gdpt_ctx *, const DScene<float> &, int, const double *, double *);
extern template int dbg_intersect<double>(gdpt_ctx *, const DScene<double> &, int, const double *, const double *, int32_t *, double *);
extern template int dbg_intersect<float>(gdpt_ctx *, const DScene<float> &, int, const double *, const double *, int32_t *, double *);
extern template int dbg_bsdf<double>(gdpt_ctx *, const DScene<double> &, int, int, const double *, const double *, const double *, double *);
extern template int dbg_bsdf<float>(gdpt_ctx *, const DScene<float> &, int, int, const double *, const double *, const double *, double *);
int fill_default_records(gdpt_ctx *ctx, int n, double *d_rec);
int pcg_fill_device(unsigned long long stream, unsigned long long seed, int n, double *d_out, unsigned int *d_raw);
} // namespace gdpt

using namespace gdpt;

namespace {

struct Resolved { int strip_count, strip_phase; int spp, sample0, y0, y1, precision, wave; unsigned long long seed; double alpha; int max_depth, rr_depth, offset_mode; };

int resolve(gdpt_ctx *c, const gdpt_render_params *p, Resolved &r) {
    if (!c || !p) return fail(GDPT_ERR_ARG, "null argument");
    if (c->bvh_pending) return fail(GDPT_ERR_ARG, "the context was created with GDPT_CREATE_NO_BVH and has not been given a BVH (gdpt_bvh_adopt)");
    r.spp = p->spp > 0 ? p->spp : c->spp;
    r.sample0 = p->sample0;
    r.y0 = p->y0 > 0 ? p->y0 : 0; r.y1 = p->y1 > 0 ? std::min(p->y1, c->height) : c->height;
    if (r.y0 >= r.y1) return fail(GDPT_ERR_ARG, "empty row band [%d,%d)", r.y0, r.y1);
    r.precision = p->precision;
    if (r.precision != GDPT_PRECISION_F64 && r.precision != GDPT_PRECISION_F32) return fail(GDPT_ERR_ARG, "unknown precision %d", r.precision);
    r.seed = p->seed ? p->seed : 0x31e241f862a1fb5eULL;
    r.alpha = p->data_cost > 0 ? p->data_cost : 0.04;
    r.max_depth = p->max_depth == -2 ? c->max_depth : p->max_depth;
    r.rr_depth = p->rr_depth == -2 ? c->rr_depth : p->rr_depth;
    r.wave = p->wave_samples;
    r.strip_count = p->strip_count > 1 ? p->strip_count : 1; r.strip_phase = p->strip_count > 1 ? p->strip_phase : 0;
    if (r.strip_phase < 0 || r.strip_phase >= r.strip_count) return fail(GDPT_ERR_ARG, "strip_phase %d outside [0, %d)", r.strip_phase, r.strip_count);
    r.offset_mode = p->offset_mode;
    if (r.offset_mode != GDPT_OFFSETS_REFERENCE && r.offset_mode != GDPT_OFFSETS_SHIFT) return fail(GDPT_ERR_ARG, "unknown offset mode %d", r.offset_mode);
    return 0;
}

template <typename R> DScene<R> &scene_of(gdpt_ctx *c);
template <> DScene<double> &scene_of<double>(gdpt_ctx *c) { return c->sc_d; }
template <> DScene<float> &scene_of<float>(gdpt_ctx *c) { return c->sc_f; }

// PATH: the plain `path` integrator (one image) instead of gradpath (five)
template <typename R, bool PATH = false>
int accumulate(gdpt_ctx *c, const Resolved &r, double *d_img, double *d_cx0, double *d_cy0, double *d_cx1, double *d_cy1) {
    DScene<R> sc = scene_of<R>(c);
    sc.max_depth = r.max_depth; sc.rr_depth = r.rr_depth;
    WaveParams wp; memset(&wp, 0, sizeof(wp));
    wp.width = c->width; wp.height = c->height; wp.y0 = r.y0; wp.y1 = r.y1;
    wp.tiles_x = (c->width + 7) / 8;
    int tiles_y = (r.y1 - r.y0 + 3) / 4;            // 4-row strips of the band ...
    wp.strip_count = r.strip_count; wp.strip_phase = r.strip_phase;
    if (r.strip_count > 1) tiles_y = tiles_y > r.strip_phase ? (tiles_y - r.strip_phase + r.strip_count - 1) / r.strip_count : 0;   // ... of which this call owns every strip_count-th
    if (tiles_y == 0) return 0;
    wp.npix_pad = (long long)wp.tiles_x * tiles_y * 32;
    wp.spp_total = r.spp; wp.seed = r.seed;
    wp.img = d_img; wp.cx0 = d_cx0; wp.cy0 = d_cy0; wp.cx1 = d_cx1; wp.cy1 = d_cy1;
    long long target = 4LL << 20;     // ~4M paths in flight per wave
    int k = r.wave > 0 ? r.wave : (int)std::max<long long>(1, target / wp.npix_pad);
    if (int rc = join_wave_streams(c)) return rc;      // (a failed earlier call may have left the second wave stream forked)
    for (int s0 = 0; s0 < r.spp; s0 += k) {
        int kk = std::min(k, r.spp - s0);
        wp.sample0 = r.sample0 + s0; wp.n_samples = kk;
        long long n = wp.npix_pad * kk;
        if (n > 0x7fffffffLL / 8) return fail(GDPT_ERR_ARG, "wave too large");
        wp.n_paths = (int)n;
        if (int rc = PATH ? run_path_wave<R>(c, sc, wp) : r.offset_mode == GDPT_OFFSETS_SHIFT ? run_shift_wave<R>(c, sc, wp) : run_wave<R>(c, sc, wp)) return rc;
    }
    return join_wave_streams(c);
}

int path_supported(gdpt_ctx *c) {
    if (c->has_envmap && !c->envmap_ready) return fail(GDPT_ERR_UNSUPPORTED, "the scene has an environment map whose image / transforms were not provided (gdpt_scene_desc.envmap_*)");
    return 0;
}

int ensure_images(gdpt_ctx *c) {
    size_t n = (size_t)c->width * c->height * 3 * sizeof(double);
    gdpt::DevBuf *bufs[] = {&c->img, &c->cx0, &c->cy0, &c->cx1, &c->cy1, &c->cx, &c->cy, &c->out};
    for (auto *b : bufs) if (int r = b->ensure(n)) return r;
    return 0;
}

} // namespace

extern "C" {

const char *gdpt_last_error(void) { return last_error().c_str(); }

int gdpt_device_count(void) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; } return n; }

int gdpt_create_ex(const gdpt_scene_desc *scene, int device, int flags, gdpt_ctx **out) {
    if (out) *out = nullptr;
    if (flags & ~(GDPT_CREATE_NO_BVH | GDPT_CREATE_BVH_HOST | GDPT_CREATE_BVH_DEVICE)) return fail(GDPT_ERR_ARG, "unknown creation flags %d", flags);
    if ((flags & GDPT_CREATE_BVH_HOST) && (flags & GDPT_CREATE_BVH_DEVICE)) return fail(GDPT_ERR_ARG, "GDPT_CREATE_BVH_HOST and GDPT_CREATE_BVH_DEVICE exclude each other");
    gdpt_ctx *c = nullptr;
    int r = create_context(scene, device, flags, &c);
    if (r != 0) { destroy_context(c); return r; }
    *out = c; return 0;
}

int gdpt_create(const gdpt_scene_desc *scene, int device, gdpt_ctx **out) { return gdpt_create_ex(scene, device, 0, out); }

int gdpt_bvh_info(gdpt_ctx *ctx, int64_t *n_nodes, int64_t *n_prims, int32_t *depth, void **d_nodes, void **d_prims) {
    if (!ctx) return fail(GDPT_ERR_ARG, "null argument");
    if (ctx->bvh_pending) return fail(GDPT_ERR_ARG, "the context has no BVH yet");
    if (n_nodes) *n_nodes = ctx->bvh.n_nodes;
    if (n_prims) *n_prims = ctx->bvh.n_prims;
    if (depth) *depth = ctx->bvh_depth;
    if (d_nodes) *d_nodes = ctx->bvh_nodes.p;
    if (d_prims) *d_prims = ctx->bvh_prims.p;
    return 0;
}

int gdpt_bvh_adopt(gdpt_ctx *ctx, const void *d_nodes, int64_t n_nodes, const void *d_prims, int64_t n_prims, int32_t depth) {
    if (!ctx || n_nodes < 0 || n_prims < 0 || (n_nodes > 0 && !d_nodes) || (n_prims > 0 && !d_prims)) return fail(GDPT_ERR_ARG, "bad argument");
    if (!ctx->bvh_pending) return fail(GDPT_ERR_ARG, "the context already has a BVH");
    if (n_nodes > 0x7fffffff || n_prims > 0x7fffffff || depth < 0) return fail(GDPT_ERR_ARG, "bad BVH size");
    if (3 * depth + 1 > STACK_TOTAL) return fail(GDPT_ERR_UNSUPPORTED, "BVH depth %d exceeds the traversal stack", depth);
    GD_CUDA(cudaSetDevice(ctx->device));
    if (int r = ctx->bvh_nodes.ensure(std::max<size_t>((size_t)n_nodes * sizeof(BvhNode), 16))) return r;
    if (int r = ctx->bvh_prims.ensure(std::max<size_t>((size_t)n_prims * sizeof(BvhPrim), 16))) return r;
    if (n_nodes) GD_CUDA(cudaMemcpyAsync(ctx->bvh_nodes.p, d_nodes, (size_t)n_nodes * sizeof(BvhNode), cudaMemcpyDefault, ctx->stream));
    if (n_prims) GD_CUDA(cudaMemcpyAsync(ctx->bvh_prims.p, d_prims, (size_t)n_prims * sizeof(BvhPrim), cudaMemcpyDefault, ctx->stream));
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->bvh.nodes = ctx->bvh_nodes.as<float4>(); ctx->bvh.prims = ctx->bvh_prims.as<float4>();
    ctx->bvh.n_nodes = (int)n_nodes; ctx->bvh.n_prims = (int)n_prims;
    ctx->bvh_depth = depth;
    ctx->bvh_pending = false;
    return 0;
}

void gdpt_destroy(gdpt_ctx *ctx) { destroy_context(ctx); }

int gdpt_host_alloc(size_t bytes, void **out) {
    if (!out || bytes == 0) return fail(GDPT_ERR_ARG, "bad argument");
    *out = nullptr;
    GD_CUDA(cudaMallocHost(out, bytes));
    return 0;
}

void gdpt_host_free(void *p) { if (p) cudaFreeHost(p); }

int gdpt_scene_bounds(gdpt_ctx *ctx, double *rc) {
    if (!ctx || !rc) return fail(GDPT_ERR_ARG, "null argument");
    for (int k = 0; k < 4; k++) rc[k] = ctx->bounds[k];
    return 0;
}

int gdpt_synchronize(gdpt_ctx *ctx) {
    if (!ctx) return fail(GDPT_ERR_ARG, "null argument");
    GD_CUDA(cudaSetDevice(ctx->device));
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int gdpt_set_stream(gdpt_ctx *ctx, void *cuda_stream) {
    if (!ctx) return fail(GDPT_ERR_ARG, "null argument");
    GD_CUDA(cudaSetDevice(ctx->device));
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    if (cuda_stream) {
        if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
        ctx->stream = (cudaStream_t)cuda_stream; ctx->own_stream = false;
    } else if (!ctx->own_stream) {
        GD_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)); ctx->own_stream = true;
    }
    return 0;
}

int gdpt_traversal_counts(gdpt_ctx *ctx, const gdpt_render_params *params, double *out3) {
    Resolved r;
    if (int rc = resolve(ctx, params, r)) return rc;
    if (!out3) return fail(GDPT_ERR_ARG, "null argument");
    GD_CUDA(cudaSetDevice(ctx->device));
    if (int rc = ensure_images(ctx)) return rc;
    if (int rc = ctx->trav_counters.ensure(64)) return rc;
    GD_CUDA(cudaMemsetAsync(ctx->trav_counters.p, 0, 64, ctx->stream));
    size_t bytes = (size_t)ctx->width * ctx->height * 3 * sizeof(double);
    gdpt::DevBuf *acc[] = {&ctx->img, &ctx->cx0, &ctx->cy0, &ctx->cx1, &ctx->cy1};
    for (auto *b : acc) GD_CUDA(cudaMemsetAsync(b->p, 0, bytes, ctx->stream));
    long long rays0 = ctx->stats.rays_traced;
    ctx->count_traversal = true;
    int rc = r.precision == GDPT_PRECISION_F64
                 ? accumulate<double>(ctx, r, ctx->img.as<double>(), ctx->cx0.as<double>(), ctx->cy0.as<double>(), ctx->cx1.as<double>(), ctx->cy1.as<double>())
                 : accumulate<float>(ctx, r, ctx->img.as<double>(), ctx->cx0.as<double>(), ctx->cy0.as<double>(), ctx->cx1.as<double>(), ctx->cy1.as<double>());
    ctx->count_traversal = false;
    if (rc) return rc;
    unsigned long long h[2];
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    GD_CUDA(cudaMemcpy(h, ctx->trav_counters.p, 16, cudaMemcpyDeviceToHost));
    double rays = (double)(ctx->stats.rays_traced - rays0);
    out3[0] = rays > 0 ? h[0] / rays : 0; out3[1] = rays > 0 ? h[1] / rays : 0; out3[2] = rays;
    return 0;
}

// L2 read-bandwidth probe: every SM streams a buffer that fits in L2 with 16-byte ld.global.cg loads.
static __global__ void __launch_bounds__(256) k_l2_read(const int4 *buf, size_t n16, int iters, int *sink) {
    int4 acc = make_int4(0, 0, 0, 0);
    for (int it = 0; it < iters; it++)
        for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
            int4 v = __ldcg(buf + i);
            acc.x ^= v.x; acc.y ^= v.y; acc.z ^= v.z; acc.w ^= v.w;
        }
    if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x5a5a5a5a) *sink = 1;   // never true for a zeroed buffer: keeps the loads alive
}

int gdpt_measure_l2_read(int device, int mbytes, int iters, double *gbs) {
    if (!gbs || mbytes < 1 || iters < 1) return fail(GDPT_ERR_ARG, "bad argument");
    GD_CUDA(cudaSetDevice(device));
    size_t bytes = (size_t)mbytes << 20, n16 = bytes / 16;
    void *buf = nullptr; int *sink = nullptr;
    GD_CUDA(cudaMalloc(&buf, bytes));
    GD_CUDA(cudaMalloc(&sink, 4));
    GD_CUDA(cudaMemset(buf, 0, bytes));
    cudaEvent_t e0, e1;
    GD_CUDA(cudaEventCreate(&e0)); GD_CUDA(cudaEventCreate(&e1));
    k_l2_read<<<148 * 8, 256>>>((const int4 *)buf, n16, 2, sink);   // warm the L2
    GD_CUDA(cudaEventRecord(e0));
    k_l2_read<<<148 * 8, 256>>>((const int4 *)buf, n16, iters, sink);
    GD_CUDA(cudaEventRecord(e1));
    GD_CUDA(cudaEventSynchronize(e1));
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    *gbs = (double)bytes * iters / (ms * 1e-3) / 1e9;
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(buf); cudaFree(sink);
    GD_CUDA(cudaGetLastError());
    return 0;
}

int gdpt_get_stats(gdpt_ctx *ctx, gdpt_stats *out) {
    if (!ctx || !out) return fail(GDPT_ERR_ARG, "null argument");
    // rays of launches whose size was only known on the device (waves launched without host read-backs)
    GD_CUDA(cudaSetDevice(ctx->device));
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    unsigned long long dev_rays = 0;
    GD_CUDA(cudaMemcpy(&dev_rays, ctx->dev_stats.p, 8, cudaMemcpyDeviceToHost));
    GD_CUDA(cudaMemset(ctx->dev_stats.p, 0, 8));
    ctx->stats.rays_traced += (long long)dev_rays;
    *out = ctx->stats; return 0;
}

int gdpt_set_option(gdpt_ctx *ctx, const char *name, int value) {
    if (!ctx || !name) return fail(GDPT_ERR_ARG, "null argument");
    const std::string n(name);
    GD_CUDA(cudaSetDevice(ctx->device));
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    if (n == "async_waves") ctx->async_waves = value != 0;
    else if (n == "wave_streams") ctx->wave_streams = std::min((int)gdpt_ctx::MAX_SLOTS, std::max(1, value));
    else if (n == "time_kernels") ctx->time_kernels = value != 0;
    else if (n == "tail_paths") ctx->tail_paths = value;
    else if (n == "poisson_fft") ctx->poisson_fft = value != 0;
    else if (n == "deterministic") ctx->deterministic = value != 0;
    else if (n == "shade_sorted") ctx->shade_sorted = value != 0;
    else if (n == "sort_rays") ctx->sort_rays = value != 0;
    else if (n == "tri_records") { ctx->use_tri_recs = value != 0; ctx->sc_f.tri_recs = (ctx->use_tri_recs && ctx->tri_recs.p) ? ctx->tri_recs.as<float>() : nullptr;
                                    ctx->sc_d.tri_recs = (ctx->use_tri_recs && ctx->tri_recs_d.p) ? ctx->tri_recs_d.as<double>() : nullptr; }
    else if (n == "trace_variant") ctx->trace_variant = value;
    else if (n == "shade_variant") ctx->shade_variant = value;
    else return fail(GDPT_ERR_ARG, "unknown option '%s'", name);
    ctx->cal.valid = false;
    return 0;
}

int gdpt_render_accumulate_device(gdpt_ctx *ctx, const gdpt_render_params *params, double *d_img, double *d_cx0, double *d_cy0,
                                  double *d_cx1, double *d_cy1) {
    Resolved r;
    if (int rc = resolve(ctx, params, r)) return rc;
    if (!d_img || !d_cx0 || !d_cy0 || !d_cx1 || !d_cy1) return fail(GDPT_ERR_ARG, "null device image");
    GD_CUDA(cudaSetDevice(ctx->device));
    return r.precision == GDPT_PRECISION_F64 ? accumulate<double>(ctx, r, d_img, d_cx0, d_cy0, d_cx1, d_cy1)
                                             : accumulate<float>(ctx, r, d_img, d_cx0, d_cy0, d_cx1, d_cy1);
}

int gdpt_merge_device(gdpt_ctx *ctx, int w, int h, const double *d_cx0, const double *d_cx1, const double *d_cy0, const double *d_cy1,
                      double *d_cx, double *d_cy) {
    if (!ctx) return fail(GDPT_ERR_ARG, "null argument");
    GD_CUDA(cudaSetDevice(ctx->device));
    return merge_device(ctx, w, h, 0, h, d_cx0, d_cx1, d_cy0, d_cy1, d_cx, d_cy);
}

int gdpt_merge_rows_device(gdpt_ctx *ctx, int w, int h, int ya, int yb, const double *d_cx0, const double *d_cx1, const double *d_cy0,
                           const double *d_cy1, double *d_cx, double *d_cy) {
    if (!ctx || ya < 0 || yb > h || w < 1) return fail(GDPT_ERR_ARG, "bad argument");
    GD_CUDA(cudaSetDevice(ctx->device));
    return merge_device(ctx, w, h, ya, yb, d_cx0, d_cx1, d_cy0, d_cy1, d_cx, d_cy);
}

int gdpt_poisson_dist_rows_forward(gdpt_ctx *ctx, const gdpt_dist_geom *geom, const double *d_u, const double *d_gx, const double *d_gy,
                                   double alpha, double *d_send, double *d_dc_partial) {
    if (!ctx || !geom || !d_u || !d_gx || !d_gy || !d_send || !d_dc_partial) return fail(GDPT_ERR_ARG, "null argument");
    GD_CUDA(cudaSetDevice(ctx->device));
    return poisson_dist_rows_forward(ctx, geom, d_u, d_gx, d_gy, alpha > 0 ? alpha : 0.04, d_send, d_dc_partial);
}

int gdpt_poisson_dist_columns(gdpt_ctx *ctx, const gdpt_dist_geom *geom, double alpha, const double *d_dc, const double *d_recv, double *d_send_back) {
    if (!ctx || !geom || !d_dc || !d_recv || !d_send_back) return fail(GDPT_ERR_ARG, "null argument");
    GD_CUDA(cudaSetDevice(ctx->device));
    return poisson_dist_columns(ctx, geom, alpha > 0 ? alpha : 0.04, d_dc, d_recv, d_send_back);
}

int gdpt_poisson_dist_rows_inverse(gdpt_ctx *ctx, const gdpt_dist_geom *geom, const double *d_recv_back, double *d_out) {
    if (!ctx || !geom || !d_recv_back || !d_out) return fail(GDPT_ERR_ARG, "null argument");
    GD_CUDA(cudaSetDevice(ctx->device));
    return poisson_dist_rows_inverse(ctx, geom, d_recv_back, d_out);
}

int gdpt_poisson_device(gdpt_ctx *ctx, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, double alpha, double *d_out) {
    if (!ctx) return fail(GDPT_ERR_ARG, "null argument");
    GD_CUDA(cudaSetDevice(ctx->device));
    return poisson_device(ctx, w, h, d_u, d_gx, d_gy, alpha > 0 ? alpha : 0.04, d_out);
}

int gdpt_render(gdpt_ctx *ctx, const gdpt_render_params *params, gdpt_images *out) {
    Resolved r;
    if (int rc = resolve(ctx, params, r)) return rc;
    GD_CUDA(cudaSetDevice(ctx->device));
    if (int rc = ensure_images(ctx)) return rc;
    const int w = ctx->width, h = ctx->height;
    size_t bytes = (size_t)w * h * 3 * sizeof(double);
    cudaStream_t st = ctx->stream;
    ctx->stats.ms_trace = ctx->stats.ms_shade = 0;
    GD_CUDA(cudaEventRecord(ctx->ev[0], st));
    gdpt::DevBuf *acc[] = {&ctx->img, &ctx->cx0, &ctx->cy0, &ctx->cx1, &ctx->cy1};
    for (auto *b : acc) GD_CUDA(cudaMemsetAsync(b->p, 0, bytes, st));
    int rc = r.precision == GDPT_PRECISION_F64
                 ? accumulate<double>(ctx, r, ctx->img.as<double>(), ctx->cx0.as<double>(), ctx->cy0.as<double>(), ctx->cx1.as<double>(), ctx->cy1.as<double>())
                 : accumulate<float>(ctx, r, ctx->img.as<double>(), ctx->cx0.as<double>(), ctx->cy0.as<double>(), ctx->cx1.as<double>(), ctx->cy1.as<double>());
    if (rc) return rc;
    GD_CUDA(cudaEventRecord(ctx->ev[1], st));
    if (int rc2 = merge_device(ctx, w, h, 0, h, ctx->cx0.as<double>(), ctx->cx1.as<double>(), ctx->cy0.as<double>(), ctx->cy1.as<double>(),
                               ctx->cx.as<double>(), ctx->cy.as<double>())) return rc2;
    if (int rc2 = poisson_device(ctx, w, h, ctx->img.as<double>(), ctx->cx.as<double>(), ctx->cy.as<double>(), r.alpha, ctx->out.as<double>())) return rc2;
    GD_CUDA(cudaEventRecord(ctx->ev[2], st));
    if (out) {
        if (out->primal) GD_CUDA(cudaMemcpyAsync(out->primal, ctx->img.p, bytes, cudaMemcpyDeviceToHost, st));
        if (out->dx) GD_CUDA(cudaMemcpyAsync(out->dx, ctx->cx.p, bytes, cudaMemcpyDeviceToHost, st));
        if (out->dy) GD_CUDA(cudaMemcpyAsync(out->dy, ctx->cy.p, bytes, cudaMemcpyDeviceToHost, st));
        if (out->poisson) GD_CUDA(cudaMemcpyAsync(out->poisson, ctx->out.p, bytes, cudaMemcpyDeviceToHost, st));
    }
    GD_CUDA(cudaEventRecord(ctx->ev[3], st));
    GD_CUDA(cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[3]); ctx->stats.ms_total = ms;
    cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]); ctx->stats.ms_poisson = ms;
    cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]); ctx->stats.ms_copy = ms;
    return 0;
}

/* path_render (render.cpp:74-117): host output */
int gdpt_render_path(gdpt_ctx *ctx, const gdpt_render_params *params, double *out_image) {
    Resolved r;
    if (int rc = resolve(ctx, params, r)) return rc;
    if (int rc = path_supported(ctx)) return rc;
    GD_CUDA(cudaSetDevice(ctx->device));
    if (int rc = ensure_images(ctx)) return rc;
    const size_t bytes = (size_t)ctx->width * ctx->height * 3 * sizeof(double);
    cudaStream_t st = ctx->stream;
    ctx->stats.ms_trace = ctx->stats.ms_shade = 0;
    GD_CUDA(cudaEventRecord(ctx->ev[0], st));
    GD_CUDA(cudaMemsetAsync(ctx->img.p, 0, bytes, st));
    int rc = r.precision == GDPT_PRECISION_F64 ? accumulate<double, true>(ctx, r, ctx->img.as<double>(), nullptr, nullptr, nullptr, nullptr)
                                               : accumulate<float, true>(ctx, r, ctx->img.as<double>(), nullptr, nullptr, nullptr, nullptr);
    if (rc) return rc;
    GD_CUDA(cudaEventRecord(ctx->ev[1], st));
    if (out_image) GD_CUDA(cudaMemcpyAsync(out_image, ctx->img.p, bytes, cudaMemcpyDeviceToHost, st));
    GD_CUDA(cudaEventRecord(ctx->ev[3], st));
    GD_CUDA(cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[3]); ctx->stats.ms_total = ms;
    cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[3]); ctx->stats.ms_copy = ms;
    ctx->stats.ms_poisson = 0;
    return 0;
}

int gdpt_render_path_accumulate_device(gdpt_ctx *ctx, const gdpt_render_params *params, double *d_img) {
    Resolved r;
    if (int rc = resolve(ctx, params, r)) return rc;
    if (int rc = path_supported(ctx)) return rc;
    if (!d_img) return fail(GDPT_ERR_ARG, "null device image");
    GD_CUDA(cudaSetDevice(ctx->device));
    return r.precision == GDPT_PRECISION_F64 ? accumulate<double, true>(ctx, r, d_img, nullptr, nullptr, nullptr, nullptr)
                                             : accumulate<float, true>(ctx, r, d_img, nullptr, nullptr, nullptr, nullptr);
}

/* path_tracing (path_tracing.h:13-348) for n explicit (x, y, rng stream) triples; rad[n][3] */
int gdpt_path_samples(gdpt_ctx *ctx, const gdpt_render_params *params, int n, const int32_t *xs, const int32_t *ys,
                      const uint64_t *streams, double *rad) {
    Resolved r;
    if (int rc = resolve(ctx, params, r)) return rc;
    if (int rc = path_supported(ctx)) return rc;
    if (!xs || !ys || !streams || !rad || n < 0) return fail(GDPT_ERR_ARG, "bad argument");
    if (n == 0) return 0;
    for (int i = 0; i < n; i++) if (xs[i] < 0 || xs[i] >= ctx->width || ys[i] < 0 || ys[i] >= ctx->height) return fail(GDPT_ERR_ARG, "pixel %d out of range", i);
    GD_CUDA(cudaSetDevice(ctx->device));
    if (int rc = ctx->xs.ensure((size_t)n * 4)) return rc;
    if (int rc = ctx->ys.ensure((size_t)n * 4)) return rc;
    if (int rc = ctx->streams.ensure((size_t)n * 8)) return rc;
    if (int rc = ctx->recout.ensure((size_t)n * 3 * 8)) return rc;
    GD_CUDA(cudaMemcpyAsync(ctx->xs.p, xs, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    GD_CUDA(cudaMemcpyAsync(ctx->ys.p, ys, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    GD_CUDA(cudaMemcpyAsync(ctx->streams.p, streams, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
    GD_CUDA(cudaMemsetAsync(ctx->recout.p, 0, (size_t)n * 3 * 8, ctx->stream));
    WaveParams wp; memset(&wp, 0, sizeof(wp));
    wp.width = ctx->width; wp.height = ctx->height; wp.y0 = 0; wp.y1 = ctx->height; wp.tiles_x = (ctx->width + 7) / 8;
    wp.n_paths = n; wp.npix_pad = n; wp.spp_total = 1; wp.seed = r.seed;
    wp.xs = ctx->xs.as<int>(); wp.ys = ctx->ys.as<int>(); wp.streams = ctx->streams.as<unsigned long long>();
    wp.rec_out = ctx->recout.as<double>();
    int rc;
    if (r.precision == GDPT_PRECISION_F64) { DScene<double> sc = ctx->sc_d; sc.max_depth = r.max_depth; sc.rr_depth = r.rr_depth; rc = run_path_wave<double>(ctx, sc, wp); }
    else { DScene<float> sc = ctx->sc_f; sc.max_depth = r.max_depth; sc.rr_depth = r.rr_depth; rc = run_path_wave<float>(ctx, sc, wp); }
    if (rc) return rc;
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    GD_CUDA(cudaMemcpy(rad, ctx->recout.p, (size_t)n * 3 * 8, cudaMemcpyDeviceToHost));
    return 0;
}

int gdpt_sample_light(gdpt_ctx *ctx, int precision, int n, const double *ref_points, const double *rnd, double *out) {
    if (!ctx || !ref_points || !rnd || !out || n < 0) return fail(GDPT_ERR_ARG, "bad argument");
    if (int rc = path_supported(ctx)) return rc;
    if (n == 0) return 0;
    GD_CUDA(cudaSetDevice(ctx->device));
    return precision == GDPT_PRECISION_F64 ? dbg_sample_light<double>(ctx, ctx->sc_d, n, ref_points, rnd, out) : dbg_sample_light<float>(ctx, ctx->sc_f, n, ref_points, rnd, out);
}

int gdpt_occluded(gdpt_ctx *ctx, int precision, int n, const double *rays, int32_t *out) {
    if (!ctx || !rays || !out || n < 0) return fail(GDPT_ERR_ARG, "bad argument");
    if (n == 0) return 0;
    GD_CUDA(cudaSetDevice(ctx->device));
    return dbg_occluded(ctx, precision == GDPT_PRECISION_F64, n, rays, out);
}

int gdpt_render_aux(gdpt_ctx *ctx, int integrator, int precision, double *out_image) {
    if (!ctx || !out_image) return fail(GDPT_ERR_ARG, "null argument");
    if (integrator < 0 || integrator > 4) return fail(GDPT_ERR_ARG, "integrator %d is not an auxiliary buffer (0 depth .. 4 mipmapLevel)", integrator);
    GD_CUDA(cudaSetDevice(ctx->device));
    return precision == GDPT_PRECISION_F64 ? dbg_aux<double>(ctx, ctx->sc_d, integrator, out_image) : dbg_aux<float>(ctx, ctx->sc_f, integrator, out_image);
}

int gdpt_grad_samples(gdpt_ctx *ctx, const gdpt_render_params *params, int n, const int32_t *xs, const int32_t *ys,
                      const uint64_t *streams, double *rec) {
    Resolved r;
    if (int rc = resolve(ctx, params, r)) return rc;
    if (!xs || !ys || !streams || !rec || n < 0) return fail(GDPT_ERR_ARG, "bad argument");
    if (n == 0) return 0;
    for (int i = 0; i < n; i++) if (xs[i] < 0 || xs[i] >= ctx->width || ys[i] < 0 || ys[i] >= ctx->height) return fail(GDPT_ERR_ARG, "pixel %d out of range", i);
    GD_CUDA(cudaSetDevice(ctx->device));
    if (int rc = ctx->xs.ensure((size_t)n * 4)) return rc;
    if (int rc = ctx->ys.ensure((size_t)n * 4)) return rc;
    if (int rc = ctx->streams.ensure((size_t)n * 8)) return rc;
    if (int rc = ctx->recout.ensure((size_t)n * 23 * 8)) return rc;
    GD_CUDA(cudaMemcpyAsync(ctx->xs.p, xs, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    GD_CUDA(cudaMemcpyAsync(ctx->ys.p, ys, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    GD_CUDA(cudaMemcpyAsync(ctx->streams.p, streams, (size_t)n * 8, cudaMemcpyHostToDevice, ctx->stream));
    if (int rc = fill_default_records(ctx, n, ctx->recout.as<double>())) return rc;
    WaveParams wp; memset(&wp, 0, sizeof(wp));
    wp.width = ctx->width; wp.height = ctx->height; wp.y0 = 0; wp.y1 = ctx->height; wp.tiles_x = (ctx->width + 7) / 8;
    wp.n_paths = n; wp.npix_pad = n; wp.spp_total = 1; wp.seed = r.seed;
    wp.xs = ctx->xs.as<int>(); wp.ys = ctx->ys.as<int>(); wp.streams = ctx->streams.as<unsigned long long>();
    wp.rec_out = ctx->recout.as<double>();
    int rc;
    const bool shift = r.offset_mode == GDPT_OFFSETS_SHIFT;
    if (r.precision == GDPT_PRECISION_F64) { DScene<double> sc = ctx->sc_d; sc.max_depth = r.max_depth; sc.rr_depth = r.rr_depth; rc = shift ? run_shift_wave<double>(ctx, sc, wp) : run_wave<double>(ctx, sc, wp); }
    else { DScene<float> sc = ctx->sc_f; sc.max_depth = r.max_depth; sc.rr_depth = r.rr_depth; rc = shift ? run_shift_wave<float>(ctx, sc, wp) : run_wave<float>(ctx, sc, wp); }
    if (rc) return rc;
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    GD_CUDA(cudaMemcpy(rec, ctx->recout.p, (size_t)n * 23 * 8, cudaMemcpyDeviceToHost));
    return 0;
}

} // extern "C"
