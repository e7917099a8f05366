// Host side of the screened-Poisson solve (kernels: gdpt_poisson.cuh).
#include "gdpt_context.cuh"
#include "gdpt_poisson.cuh"
#include "gdpt_poisson_fft.cuh"

namespace gdpt {

constexpr int SM_COUNT_P = 148;
static inline int blocks_for_p(long long n, int block) { return (int)((n + block - 1) / block); }
#define SM_COUNT SM_COUNT_P
#define blocks_for blocks_for_p

static int ensure_matrices(gdpt_ctx *c, int w, int h) {
    PoissonCache &pc = c->pc;
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

static int ensure_poisson(gdpt_ctx *c, int w, int h) {
    PoissonCache &pc = c->pc;
    size_t plane = (size_t)w * h * 3 * sizeof(double);
    if (int r = pc.work.ensure(plane)) return r;
    if (int r = pc.work2.ensure(plane)) return r;
    return ensure_matrices(c, w, h);
}

// ---------------------------------------------------------------- FFT form (gdpt_poisson_fft.cuh)
constexpr int FFT_MIN_N = 16, FFT_MAX_P = 8192;      // P complex + N real of shared memory per block: 160 KB at P = 8192 (N <= 4097)
static int ilog2(int P) { int l = 0; while ((1 << l) < P) l++; return l; }
// P complex (the transform) + the pass twiddles + N real (the line) of shared memory per block
static size_t fft_smem(int P, int N) { return ((size_t)P + fft_twiddle_count(P, ilog2(P))) * sizeof(double2) + (size_t)((N + 1) & ~1) * sizeof(double); }
static bool fft_eligible(int N) {
    if (N < FFT_MIN_N) return false;
    int P = 16; while (P < 2 * (N - 1) - 1) P <<= 1;
    return P <= FFT_MAX_P;
}
static int ensure_plan(gdpt_ctx *c, DctPlanBufs &b, int N) {
    if (b.N == N) return 0;
    const int M = N - 1;
    int P = 16, logP = 4; while (P < 2 * M - 1) { P <<= 1; logP++; }
    if (int r = b.tw.ensure((size_t)std::max(1, fft_twiddle_count(P, logP)) * sizeof(double2))) return r;
    if (int r = b.chirp.ensure((size_t)M * sizeof(double2))) return r;
    if (int r = b.wpost.ensure((size_t)M * sizeof(double2))) return r;
    if (int r = b.bhat.ensure((size_t)P * sizeof(double2))) return r;
    static bool attr_set = false;
    if (!attr_set) {
        GD_CUDA(cudaFuncSetAttribute(k_dct1_fft, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fft_smem(FFT_MAX_P, FFT_MAX_P / 2 + 2)));
        GD_CUDA(cudaFuncSetAttribute(k_dct_plan_filter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(FFT_MAX_P * sizeof(double2))));
        attr_set = true;
    }
    k_dct_plan_tables<<<blocks_for(P, 256), 256, 0, c->stream>>>(M, P, logP, b.tw.as<double2>(), b.chirp.as<double2>(), b.wpost.as<double2>());
    k_dct_plan_filter<<<1, FFT_THREADS, (size_t)P * sizeof(double2), c->stream>>>(M, P, logP, b.tw.as<double2>(), b.chirp.as<double2>(), b.bhat.as<double2>());
    c->stats.kernel_launches += 2;
    GD_CUDA(cudaGetLastError());
    b.N = N; b.M = M; b.P = P; b.logP = logP;
    return 0;
}
static DctPlan plan_of(const DctPlanBufs &b) {
    DctPlan p; p.N = b.N; p.M = b.M; p.P = b.P; p.logP = b.logP;
    p.tw = b.tw.as<double2>(); p.chirp = b.chirp.as<double2>(); p.bhat = b.bhat.as<double2>(); p.wpost = b.wpost.as<double2>();
    return p;
}

// rhs -> rows (DCT-I along x) -> columns (DCT-I along y, spectral divide + DC, DCT-I along y) -> rows (DCT-I along x, scale,
// RGB interleave): 1 + 3 launches, two planar work images, the column pass in place.
static int poisson_device_fft(gdpt_ctx *c, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, double alpha, double *d_out) {
    PoissonCache &pc = c->pc;
    const size_t plane3 = (size_t)w * h * 3 * sizeof(double);
    if (int r = pc.work.ensure(plane3)) return r;
    if (int r = pc.work2.ensure(plane3)) return r;
    if (int r = pc.partial.ensure((size_t)SM_COUNT * 8 * 3 * sizeof(double))) return r;
    if (int r = pc.dc.ensure(64)) return r;
    if (int r = ensure_plan(c, pc.plan_w, w)) return r;
    if (int r = ensure_plan(c, pc.plan_h, h)) return r;
    cudaStream_t st = c->stream;
    double *A = pc.work.as<double>(), *B = pc.work2.as<double>();
    const int nb = SM_COUNT * 8;
    k_poisson_rhs<<<nb, RHS_BLOCK, 0, st>>>(w, h, 0, h, d_u, d_gx, d_gy, alpha, A, pc.partial.as<double>());
    k_dc_reduce<<<1, 96, 0, st>>>(nb, pc.partial.as<double>(), pc.dc.as<double>());
    const DctPlan pw = plan_of(pc.plan_w), ph = plan_of(pc.plan_h);
    DctArgs a; memset(&a, 0, sizeof(a));
    a.plane_stride = (long long)w * h; a.img_w = w; a.img_h = h; a.alpha = alpha; a.dc = pc.dc.as<double>();
    a.scale = 4.0 * (w - 1) * (h - 1); a.out = d_out;
    // rows, forward
    a.mode = DCT_STORE; a.count = 3 * h; a.per_plane = h; a.line_stride = w; a.elem_stride = 1;
    const int grid_cap = SM_COUNT * 2;      // persistent blocks: the pass twiddles are staged once per block, the lines are looped over
    k_dct1_fft<<<std::min(a.count, grid_cap), FFT_THREADS, fft_smem(pw.P, w), st>>>(pw, a, A, B);
    // columns: forward, divide, inverse — in place
    a.mode = DCT_COLUMNS; a.count = 3 * w; a.per_plane = w; a.line_stride = 1; a.elem_stride = w;
    k_dct1_fft<<<std::min(a.count, grid_cap), FFT_THREADS, fft_smem(ph.P, h), st>>>(ph, a, B, B);
    // rows, inverse + scale + interleave
    a.mode = DCT_SCALE_INTERLEAVE; a.count = 3 * h; a.per_plane = h; a.line_stride = w; a.elem_stride = 1;
    k_dct1_fft<<<std::min(a.count, grid_cap), FFT_THREADS, fft_smem(pw.P, w), st>>>(pw, a, B, nullptr);
    c->stats.kernel_launches += 5;
    GD_CUDA(cudaGetLastError());
    return 0;
}

int poisson_device(gdpt_ctx *c, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, double alpha, double *d_out) {
    if (w < 2 || h < 2) return fail(GDPT_ERR_ARG, "Poisson solve needs at least a 2x2 image");
    if (c->poisson_fft && fft_eligible(w) && fft_eligible(h)) return poisson_device_fft(c, w, h, d_u, d_gx, d_gy, alpha, d_out);
    if (int r = ensure_poisson(c, w, h)) return r;
    PoissonCache &pc = c->pc;
    cudaStream_t st = c->stream;
    double *A = pc.work.as<double>(), *B = pc.work2.as<double>();
    const int nb = SM_COUNT * 8;
    k_poisson_rhs<<<nb, RHS_BLOCK, 0, st>>>(w, h, 0, h, d_u, d_gx, d_gy, alpha, A, pc.partial.as<double>());
    k_dc_reduce<<<1, 96, 0, st>>>(nb, pc.partial.as<double>(), pc.dc.as<double>());
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


// ---------------------------------------------------------------- distributed solve (row bands <-> column bands)
// The separable transform F = D_H * X * D_W is split at the point where the data has to change owners:
//   rows_forward   (row band, local)    rhs, DC partial sums, X*D_W with the result PACKED per destination rank
//   -- all-to-all transpose (NCCL, run by the caller) --
//   columns        (column band, local) D_H*(.), spectral divide + DC, D_H*(.) again (inverse along columns)
//   -- all-to-all transpose back --
//   rows_inverse   (row band, local)    (.)*D_W read straight from the received chunks, scale, RGB interleave
// Only two transposes are needed because the DCT is applied as a dense matrix product: nothing is ever
// physically transposed, the messages are just the [rows_r x cols_s] blocks.
static int dist_check(const gdpt_dist_geom *g) {
    if (!g || g->w < 2 || g->h < 2 || g->n_ranks < 1 || g->rank < 0 || g->rank >= g->n_ranks || g->y0 < 0 || g->y1 > g->h || g->y1 < g->y0 ||
        g->band_w < 1 || (long long)g->band_w * (g->n_ranks - 1) >= g->w)
        return fail(GDPT_ERR_ARG, "bad distributed-solve geometry");
    return 0;
}
static Banded dist_band(const gdpt_dist_geom *g) { Banded b; b.n_bands = g->n_ranks; b.band_w = g->band_w; b.rows = g->y1 - g->y0; b.W = g->w; return b; }

int poisson_dist_rows_forward(gdpt_ctx *c, const gdpt_dist_geom *g, const double *d_u, const double *d_gx, const double *d_gy, double alpha,
                              double *d_send, double *d_dc_partial) {
    if (int r = dist_check(g)) return r;
    if (int r = ensure_matrices(c, g->w, g->h)) return r;
    PoissonCache &pc = c->pc;
    cudaStream_t st = c->stream;
    const int w = g->w, rows = g->y1 - g->y0, nb = SM_COUNT * 8;
    if (int r = pc.work.ensure(std::max<size_t>((size_t)w * rows * 3 * sizeof(double), 16))) return r;
    k_poisson_rhs<<<nb, RHS_BLOCK, 0, st>>>(w, g->h, g->y0, g->y1, d_u, d_gx, d_gy, alpha, pc.work.as<double>(), pc.partial.as<double>());
    k_dc_reduce<<<1, 96, 0, st>>>(nb, pc.partial.as<double>(), d_dc_partial);
    c->stats.kernel_launches += 2;
    if (rows > 0) {
        Epilogue ep; memset(&ep, 0, sizeof(ep));
        ep.mode = EPI_PACK; ep.band = dist_band(g);
        dim3 grid(blocks_for(w, GT), blocks_for(3 * rows, GT), 1);
        k_dgemm<<<grid, 256, 0, st>>>(3 * rows, w, w, pc.work.as<double>(), w, 0, pc.Dw.as<double>(), w, 0, d_send, w, 0, ep);
        c->stats.kernel_launches++;
    }
    GD_CUDA(cudaGetLastError());
    return 0;
}

int poisson_dist_columns(gdpt_ctx *c, const gdpt_dist_geom *g, double alpha, const double *d_dc, const double *d_recv, double *d_send_back) {
    if (int r = dist_check(g)) return r;
    if (int r = ensure_matrices(c, g->w, g->h)) return r;
    PoissonCache &pc = c->pc;
    cudaStream_t st = c->stream;
    const int h = g->h, x0 = g->rank * g->band_w, ws = (g->rank == g->n_ranks - 1) ? g->w - x0 : g->band_w, n = 3 * ws;
    if (int r = pc.work2.ensure((size_t)h * n * sizeof(double))) return r;
    Epilogue ep; memset(&ep, 0, sizeof(ep));
    ep.img_w = g->w; ep.img_h = h; ep.alpha = alpha; ep.dc = d_dc; ep.ch_w = ws; ep.x0 = x0;
    dim3 grid(blocks_for(n, GT), blocks_for(h, GT), 1);
    ep.mode = EPI_DIVIDE;
    k_dgemm<<<grid, 256, 0, st>>>(h, n, h, pc.DhT.as<double>(), h, 0, d_recv, n, 0, pc.work2.as<double>(), n, 0, ep);
    ep.mode = EPI_STORE;
    k_dgemm<<<grid, 256, 0, st>>>(h, n, h, pc.DhT.as<double>(), h, 0, pc.work2.as<double>(), n, 0, d_send_back, n, 0, ep);
    c->stats.kernel_launches += 2;
    GD_CUDA(cudaGetLastError());
    return 0;
}

int poisson_dist_rows_inverse(gdpt_ctx *c, const gdpt_dist_geom *g, const double *d_recv_back, double *d_out) {
    if (int r = dist_check(g)) return r;
    if (int r = ensure_matrices(c, g->w, g->h)) return r;
    PoissonCache &pc = c->pc;
    const int w = g->w, rows = g->y1 - g->y0;
    if (rows == 0) return 0;
    Epilogue ep; memset(&ep, 0, sizeof(ep));
    ep.mode = EPI_SCALE_INTERLEAVE; ep.img_w = w; ep.img_h = rows; ep.band = dist_band(g);
    ep.scale = 4.0 * (w - 1) * (g->h - 1);
    ep.out = d_out + (size_t)g->y0 * w * 3;
    dim3 grid(blocks_for(w, GT), blocks_for(3 * rows, GT), 1);
    k_dgemm<true><<<grid, 256, 0, c->stream>>>(3 * rows, w, w, d_recv_back, w, 0, pc.Dw.as<double>(), w, 0, nullptr, w, 0, ep);
    c->stats.kernel_launches++;
    GD_CUDA(cudaGetLastError());
    return 0;
}

int merge_device(gdpt_ctx *ctx, int w, int h, int ya, int yb, const double *d_cx0, const double *d_cx1, const double *d_cy0, const double *d_cy1, double *d_cx, double *d_cy) {
    if (yb <= ya) return 0;
    k_merge<<<SM_COUNT * 8, 256, 0, ctx->stream>>>(w, h, ya, yb, d_cx0, d_cx1, d_cy0, d_cy1, d_cx, d_cy);
    ctx->stats.kernel_launches++;
    GD_CUDA(cudaGetLastError());
    return 0;
}

} // namespace gdpt
