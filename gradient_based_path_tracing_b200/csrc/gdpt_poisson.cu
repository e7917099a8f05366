// Host side of the screened-Poisson solve (kernels: gdpt_poisson.cuh).
#include "gdpt_context.cuh"
#include "gdpt_poisson.cuh"

namespace gdpt {

constexpr int SM_COUNT_P = 148;
static inline int blocks_for_p(long long n, int block) { return (int)((n + block - 1) / block); }
#define SM_COUNT SM_COUNT_P
#define blocks_for blocks_for_p

static int ensure_poisson(gdpt_ctx *c, int w, int h) {
    PoissonCache &pc = c->pc;
    size_t plane = (size_t)w * h * 3 * sizeof(double);
    if (int r = pc.work.ensure(plane)) return r;
    if (int r = pc.work2.ensure(plane)) return r;
    if (int r = pc.partial.ensure((size_t)SM_COUNT * 8 * 3 * sizeof(double))) return r;
    if (int r = pc.dc.ensure(64)) return r;
    if (pc.w != w || pc.h != h) {
        if (int r = pc.Dw.ensure((size_t)w * w * sizeof(double))) return r;
        if (int r = pc.DhT.ensure((size_t)h * h * sizeof(double))) return r;
        k_dct_matrix<<<SM_COUNT * 4, 256, 0, c->stream>>>(w, 0, pc.Dw.as<double>());
        k_dct_matrix<<<SM_COUNT * 4, 256, 0, c->stream>>>(h, 1, pc.DhT.as<double>());
        c->stats.kernel_launches += 2;
        pc.w = w; pc.h = h;
    }
    return 0;
}

int poisson_device(gdpt_ctx *c, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, double alpha, double *d_out) {
    if (w < 2 || h < 2) return fail(GDPT_ERR_ARG, "Poisson solve needs at least a 2x2 image");
    if (int r = ensure_poisson(c, w, h)) return r;
    PoissonCache &pc = c->pc;
    cudaStream_t st = c->stream;
    double *A = pc.work.as<double>(), *B = pc.work2.as<double>();
    const int nb = SM_COUNT * 8;
    k_poisson_rhs<<<nb, RHS_BLOCK, 0, st>>>(w, h, d_u, d_gx, d_gy, alpha, A, pc.partial.as<double>());
    k_dc_reduce<<<1, 32, 0, st>>>(nb, pc.partial.as<double>(), pc.dc.as<double>());
    Epilogue ep; memset(&ep, 0, sizeof(ep));
    ep.img_w = w; ep.img_h = h; ep.alpha = alpha; ep.dc = pc.dc.as<double>(); ep.out = d_out;
    ep.scale = 4.0 * (w - 1) * (h - 1);
    long long plane = (long long)w * h;
    dim3 g_rows(blocks_for(w, GT), blocks_for(3 * h, GT), 1), g_cols(blocks_for(w, GT), blocks_for(h, GT), 3);
    // forward: rows then columns (+ spectral divide, DC)
    ep.mode = EPI_STORE;
    k_dgemm<<<g_rows, 256, 0, st>>>(3 * h, w, w, A, w, 0, pc.Dw.as<double>(), w, 0, B, w, 0, ep);
    ep.mode = EPI_DIVIDE;
    k_dgemm<<<g_cols, 256, 0, st>>>(h, w, h, pc.DhT.as<double>(), h, 0, B, w, plane, A, w, plane, ep);
    // inverse (DCT-I is its own inverse up to the scale): columns then rows (+ scale, re-interleave)
    ep.mode = EPI_STORE;
    k_dgemm<<<g_cols, 256, 0, st>>>(h, w, h, pc.DhT.as<double>(), h, 0, A, w, plane, B, w, plane, ep);
    ep.mode = EPI_SCALE_INTERLEAVE;
    k_dgemm<<<g_rows, 256, 0, st>>>(3 * h, w, w, B, w, 0, pc.Dw.as<double>(), w, 0, A, w, 0, ep);
    c->stats.kernel_launches += 6;
    GD_CUDA(cudaGetLastError());
    return 0;
}


int merge_device(gdpt_ctx *ctx, int w, int h, const double *d_cx0, const double *d_cx1, const double *d_cy0, const double *d_cy1, double *d_cx, double *d_cy) {
    k_merge<<<SM_COUNT * 8, 256, 0, ctx->stream>>>(w, h, d_cx0, d_cx1, d_cy0, d_cy1, d_cx, d_cy);
    ctx->stats.kernel_launches++;
    GD_CUDA(cudaGetLastError());
    return 0;
}

} // namespace gdpt
