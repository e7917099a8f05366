// C ABI, second half: the stand-alone Poisson entry point with host buffers and the
// validation entry points (gdpt_sample_primary / gdpt_intersect / gdpt_bsdf / gdpt_pcg_fill).
// Host glue only; kernels live in gdpt_validate_*.cu / gdpt_poisson.cu.
#include "gdpt_context.cuh"

namespace gdpt {
template <typename R> int dbg_sample_primary(gdpt_ctx *ctx, const DScene<R> &sc, int n, const double *screen_pos, double *rays);
template <typename R> int dbg_intersect(gdpt_ctx *ctx, const DScene<R> &sc, int n, const double *rays, const double *rdiff, int32_t *ids, double *vtx);
template <typename R> int dbg_bsdf(gdpt_ctx *ctx, const DScene<R> &sc, int mat_id, int n, const double *vtx, const double *dirs, const double *rnd, double *out);
extern template int dbg_sample_primary<double>(gdpt_ctx *, const DScene<double> &, int, const double *, double *);
extern template int dbg_sample_primary<float>(gdpt_ctx *, const DScene<float> &, int, const double *, double *);
extern template int dbg_intersect<double>(gdpt_ctx *, const DScene<double> &, int, const double *, const double *, int32_t *, double *);
extern template int dbg_intersect<float>(gdpt_ctx *, const DScene<float> &, int, const double *, const double *, int32_t *, double *);
extern template int dbg_bsdf<double>(gdpt_ctx *, const DScene<double> &, int, int, const double *, const double *, const double *, double *);
extern template int dbg_bsdf<float>(gdpt_ctx *, const DScene<float> &, int, int, const double *, const double *, const double *, double *);
int pcg_fill_device(unsigned long long stream, unsigned long long seed, int n, double *d_out, unsigned int *d_raw);
}
using namespace gdpt;

namespace {
struct Scratch {   // device scratch that frees itself
    std::vector<void *> ptrs;
    ~Scratch() { for (void *p : ptrs) cudaFree(p); }
    template <typename T> int alloc(T **out, size_t n) {
        void *p = nullptr;
        cudaError_t e = cudaMalloc(&p, std::max<size_t>(n * sizeof(T), 16));
        if (e != cudaSuccess) return fail(GDPT_ERR_CUDA, "cudaMalloc failed: %s", cudaGetErrorString(e));
        ptrs.push_back(p); *out = (T *)p; return 0;
    }
    template <typename T> int put(T **out, const T *host, size_t n) {
        if (int r = alloc(out, n)) return r;
        if (host && n) { GD_CUDA(cudaMemcpy(*out, host, n * sizeof(T), cudaMemcpyHostToDevice)); GD_CUDA(cudaStreamSynchronize((cudaStream_t)0)); }
        return 0;
    }
};
} // namespace

// implemented in gdpt_api.cu
extern "C" int gdpt_poisson_device(gdpt_ctx *ctx, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, double alpha, double *d_out);

extern "C" {

int gdpt_poisson(int device, int w, int h, const double *u, const double *gx, const double *gy, double alpha, double *out) {
    if (!u || !gx || !gy || !out) return fail(GDPT_ERR_ARG, "null argument");
    if (w < 2 || h < 2) return fail(GDPT_ERR_ARG, "Poisson solve needs at least a 2x2 image");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(GDPT_ERR_CUDA, "no CUDA device: libgdpt has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(GDPT_ERR_ARG, "device out of range");
    GD_CUDA(cudaSetDevice(device));
    // a bare context (no scene) that only carries the stream and the Poisson cache
    gdpt_ctx *c = new gdpt_ctx();
    c->device = device; memset(&c->stats, 0, sizeof(c->stats));
    for (auto &e : c->ev) e = nullptr;
    int rc = 0;
    size_t n = (size_t)w * h * 3;
    Scratch s; double *du, *dgx, *dgy, *dout;
    do {
        if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { rc = fail(GDPT_ERR_CUDA, "stream creation failed"); break; }
        if ((rc = s.put(&du, u, n))) break;
        if ((rc = s.put(&dgx, gx, n))) break;
        if ((rc = s.put(&dgy, gy, n))) break;
        if ((rc = s.alloc(&dout, n))) break;
        if ((rc = gdpt_poisson_device(c, w, h, du, dgx, dgy, alpha, dout))) break;
        if (cudaStreamSynchronize(c->stream) != cudaSuccess) { rc = fail(GDPT_ERR_CUDA, "Poisson kernels failed: %s", cudaGetErrorString(cudaGetLastError())); break; }
        if (cudaMemcpy(out, dout, n * sizeof(double), cudaMemcpyDeviceToHost) != cudaSuccess) { rc = fail(GDPT_ERR_CUDA, "copy back failed"); break; }
    } while (false);
    destroy_context(c);
    return rc;
}

int gdpt_sample_primary(gdpt_ctx *ctx, int precision, int n, const double *screen_pos, double *rays) {
    if (!ctx || !screen_pos || !rays || n < 0) return fail(GDPT_ERR_ARG, "bad argument");
    if (n == 0) return 0;
    GD_CUDA(cudaSetDevice(ctx->device));
    return precision == GDPT_PRECISION_F64 ? dbg_sample_primary<double>(ctx, ctx->sc_d, n, screen_pos, rays)
                                           : dbg_sample_primary<float>(ctx, ctx->sc_f, n, screen_pos, rays);
}

int gdpt_intersect(gdpt_ctx *ctx, int precision, int n, const double *rays, const double *rdiff, int32_t *ids, double *vtx) {
    if (!ctx || !rays || !ids || !vtx || n < 0) return fail(GDPT_ERR_ARG, "bad argument");
    if (n == 0) return 0;
    GD_CUDA(cudaSetDevice(ctx->device));
    return precision == GDPT_PRECISION_F64 ? dbg_intersect<double>(ctx, ctx->sc_d, n, rays, rdiff, ids, vtx)
                                           : dbg_intersect<float>(ctx, ctx->sc_f, n, rays, rdiff, ids, vtx);
}

int gdpt_bsdf(gdpt_ctx *ctx, int precision, int mat_id, int n, const double *vtx, const double *dirs, const double *rnd, double *out) {
    if (!ctx || !vtx || !dirs || !rnd || !out || n < 0) return fail(GDPT_ERR_ARG, "bad argument");
    if (mat_id < 0 || mat_id >= ctx->sc_d.n_materials) return fail(GDPT_ERR_ARG, "material id out of range");
    if (n == 0) return 0;
    GD_CUDA(cudaSetDevice(ctx->device));
    return precision == GDPT_PRECISION_F64 ? dbg_bsdf<double>(ctx, ctx->sc_d, mat_id, n, vtx, dirs, rnd, out)
                                           : dbg_bsdf<float>(ctx, ctx->sc_f, mat_id, n, vtx, dirs, rnd, out);
}

int gdpt_pcg_fill(int device, uint64_t stream, uint64_t seed, int n, double *out, uint32_t *raw) {
    if (!out || n < 0) return fail(GDPT_ERR_ARG, "bad argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(GDPT_ERR_CUDA, "no CUDA device: libgdpt has no CPU fallback");
    if (n == 0) return 0;
    GD_CUDA(cudaSetDevice(device));
    Scratch s; double *dout; unsigned int *draw = nullptr;
    if (int r = s.alloc(&dout, (size_t)n)) return r;
    if (raw) if (int r = s.alloc(&draw, (size_t)n)) return r;
    if (int r = pcg_fill_device(stream, seed ? seed : 0x31e241f862a1fb5eULL, n, dout, draw)) return r;
    GD_CUDA(cudaMemcpy(out, dout, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost));
    if (raw) GD_CUDA(cudaMemcpy(raw, draw, (size_t)n * sizeof(unsigned int), cudaMemcpyDeviceToHost));
    return 0;
}

} // extern "C"
