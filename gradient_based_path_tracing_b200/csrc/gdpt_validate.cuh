// Host launchers of the validation kernels (gdpt_debug.cuh), templated on the arithmetic
// type; instantiated in gdpt_validate_f64.cu / gdpt_validate_f32.cu.
#pragma once
#include "gdpt_wave.cuh"
#include "gdpt_debug.cuh"

namespace gdpt {

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

template <typename R> int dbg_sample_primary(gdpt_ctx *ctx, const DScene<R> &sc, int n, const double *screen_pos, double *rays) {
    Scratch s; double *dsp, *dr;
    if (int r = s.put(&dsp, screen_pos, (size_t)n * 2)) return r;
    if (int r = s.alloc(&dr, (size_t)n * 6)) return r;
    k_dbg_sample_primary<R><<<blocks_for(n, 128), 128, 0, ctx->stream>>>(sc, n, dsp, dr);
    ctx->stats.kernel_launches++;
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    GD_CUDA(cudaMemcpy(rays, dr, (size_t)n * 6 * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}

template <typename R> int dbg_intersect(gdpt_ctx *ctx, const DScene<R> &sc, int n, const double *rays, const double *rdiff, int32_t *ids, double *vtx) {
    Scratch s; double *dr, *drd = nullptr, *dv; int *di;
    if (int r = s.put(&dr, rays, (size_t)n * 8)) return r;
    if (rdiff) if (int r = s.put(&drd, rdiff, (size_t)n * 2)) return r;
    if (int r = s.alloc(&di, (size_t)n * 3)) return r;
    if (int r = s.alloc(&dv, (size_t)n * 24)) return r;
    k_dbg_intersect<R, ModeSel<R>::exact><<<blocks_for(n, TRACE_BLOCK), TRACE_BLOCK, 0, ctx->stream>>>(sc, ctx->bvh, n, dr, drd, di, dv);
    ctx->stats.kernel_launches++;
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    GD_CUDA(cudaMemcpy(ids, di, (size_t)n * 3 * sizeof(int), cudaMemcpyDeviceToHost));
    GD_CUDA(cudaMemcpy(vtx, dv, (size_t)n * 24 * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}

template <typename R> int dbg_bsdf(gdpt_ctx *ctx, const DScene<R> &sc, int mat_id, int n, const double *vtx, const double *dirs, const double *rnd, double *out) {
    Scratch s; double *dv, *dd, *drn, *dout;
    if (int r = s.put(&dv, vtx, (size_t)n * 15)) return r;
    if (int r = s.put(&dd, dirs, (size_t)n * 6)) return r;
    if (int r = s.put(&drn, rnd, (size_t)n * 3)) return r;
    if (int r = s.alloc(&dout, (size_t)n * 10)) return r;
    k_dbg_bsdf<R><<<blocks_for(n, 64), 64, 0, ctx->stream>>>(sc, mat_id, n, dv, dd, drn, dout);
    ctx->stats.kernel_launches++;
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    GD_CUDA(cudaMemcpy(out, dout, (size_t)n * 10 * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}

template <typename R> int dbg_sample_light(gdpt_ctx *ctx, const DScene<R> &sc, int n, const double *ref_points, const double *rnd, double *out) {
    if (sc.n_lights <= 0) return fail(GDPT_ERR_ARG, "the scene has no lights");
    Scratch s; double *dp, *drn, *dout;
    if (int r = s.put(&dp, ref_points, (size_t)n * 3)) return r;
    if (int r = s.put(&drn, rnd, (size_t)n * 4)) return r;
    if (int r = s.alloc(&dout, (size_t)n * 9)) return r;
    k_dbg_sample_light<R><<<blocks_for(n, 128), 128, 0, ctx->stream>>>(sc, n, dp, drn, dout);
    ctx->stats.kernel_launches++;
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    GD_CUDA(cudaMemcpy(out, dout, (size_t)n * 9 * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}

template <typename R> int dbg_aux(gdpt_ctx *ctx, const DScene<R> &sc, int kind, double *out) {
    Scratch s; double *dimg;
    const size_t n = (size_t)sc.width * sc.height;
    if (int r = s.alloc(&dimg, n * 3)) return r;
    k_aux<R, ModeSel<R>::exact><<<blocks_for((long long)n, TRACE_BLOCK), TRACE_BLOCK, 0, ctx->stream>>>(sc, ctx->bvh, kind, dimg);
    ctx->stats.kernel_launches++;
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    GD_CUDA(cudaMemcpy(out, dimg, n * 3 * sizeof(double), cudaMemcpyDeviceToHost));
    return 0;
}

} // namespace gdpt
