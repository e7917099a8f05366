// validation launchers, R = double: explicit instantiation
#include "gdpt_validate.cuh"
namespace gdpt {
template int dbg_sample_primary<double>(gdpt_ctx *, const DScene<double> &, int, const double *, double *);
template int dbg_intersect<double>(gdpt_ctx *, const DScene<double> &, int, const double *, const double *, int32_t *, double *);
template int dbg_bsdf<double>(gdpt_ctx *, const DScene<double> &, int, int, const double *, const double *, const double *, double *);
template int dbg_sample_light<double>(gdpt_ctx *, const DScene<double> &, int, const double *, const double *, double *);
template int dbg_aux<double>(gdpt_ctx *, const DScene<double> &, int, double *);

// occluded (intersection.cpp:66-85) through the renderer's any-hit traversal kernel: rays[n][8] -> out[n] (1 = occluded)
static __global__ void k_dbg_pack_rays(int n, const double *rays, RayF *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double *q = rays + 8 * i;
    RayF r; r.ox = (float)q[0]; r.oy = (float)q[1]; r.oz = (float)q[2]; r.tnear = (float)q[6]; r.dx = (float)q[3]; r.dy = (float)q[4]; r.dz = (float)q[5]; r.tfar = (float)q[7];
    out[i] = r;
}
static __global__ void k_dbg_unpack_occluded(int n, const HitRec *hits, int *out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = hits[i].shape_id >= 0 ? 1 : 0;
}
int dbg_occluded(gdpt_ctx *ctx, bool exact, int n, const double *rays, int32_t *out) {
    Scratch s; double *dr; RayF *drays; HitRec *dh; int *dout;
    if (int r = s.put(&dr, rays, (size_t)n * 8)) return r;
    if (int r = s.alloc(&drays, (size_t)n)) return r;
    if (int r = s.alloc(&dh, (size_t)n)) return r;
    if (int r = s.alloc(&dout, (size_t)n)) return r;
    k_dbg_pack_rays<<<blocks_for(n, 128), 128, 0, ctx->stream>>>(n, dr, drays);
    int grid = std::min(blocks_for(n, TRACE_BLOCK), SM_COUNT * 16);
    if (exact) k_trace<true, false, TRACE_DEFAULT_VARIANT, true><<<grid, TRACE_BLOCK, 0, ctx->stream>>>(ctx->bvh, drays, dh, nullptr, n, nullptr);
    else k_trace<false, false, TRACE_DEFAULT_VARIANT, true><<<grid, TRACE_BLOCK, 0, ctx->stream>>>(ctx->bvh, drays, dh, nullptr, n, nullptr);
    k_dbg_unpack_occluded<<<blocks_for(n, 128), 128, 0, ctx->stream>>>(n, dh, dout);
    ctx->stats.kernel_launches += 3;
    GD_CUDA(cudaStreamSynchronize(ctx->stream));
    GD_CUDA(cudaMemcpy(out, dout, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost));
    return 0;
}

// small non-template launchers
int fill_default_records(gdpt_ctx *ctx, int n, double *d_rec) {
    k_fill_default_records<<<blocks_for(n, 128), 128, 0, ctx->stream>>>(n, d_rec);
    ctx->stats.kernel_launches++;
    GD_CUDA(cudaGetLastError());
    return 0;
}
int pcg_fill_device(unsigned long long stream, unsigned long long seed, int n, double *d_out, unsigned int *d_raw) {
    k_dbg_pcg<<<1, 32>>>(stream, seed, n, d_out, d_raw);
    GD_CUDA(cudaDeviceSynchronize());
    return 0;
}
}
