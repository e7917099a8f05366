// validation launchers, R = double: explicit instantiation
#include "gdpt_validate.cuh"
namespace gdpt {
template int dbg_sample_primary<double>(gdpt_ctx *, const DScene<double> &, int, const double *, double *);
template int dbg_intersect<double>(gdpt_ctx *, const DScene<double> &, int, const double *, const double *, int32_t *, double *);
template int dbg_bsdf<double>(gdpt_ctx *, const DScene<double> &, int, int, const double *, const double *, const double *, double *);

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
