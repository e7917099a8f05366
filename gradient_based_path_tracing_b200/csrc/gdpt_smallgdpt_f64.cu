// gdpt_smallgdpt_render — the reference's stand-alone program small_gdpt.cpp (main(), :439-563) on the GPU: its sphere
// scene rendered with the arithmetic of gdpt_smallgdpt.cuh, then the product's own merge (k_merge) and screened-Poisson
// solve (the FFT form of gdpt_poisson.cu) in place of :530-541.  Compiled with -fmad=false (build.sh, *_f64.cu) like the
// other exact-mode units: the reference is gcc on x86-64 without FMA contraction.
//
// Two stream layouts:
//   GDPT_SMALLGDPT_ROWS    the reference's: ONE erand48 stream per film row, consumed pixel after pixel, sample after sample
//                          (:462) — how many numbers a pixel draws depends on its paths, so a row cannot be split.  One warp
//                          per row: lane 0 draws the base path into shared memory (the only consumer of random numbers),
//                          lanes 0..3 then walk the four offset paths of that base path side by side.  height warps in all.
//   GDPT_SMALLGDPT_PIXELS  a stream per pixel (splitmix64 of the pixel index): every pixel a thread, 8x4 pixel tiles per
//                          warp.  Not bitwise the reference's images — the same estimator on other random numbers.
#include "gdpt_context.cuh"
#include "gdpt_smallgdpt.cuh"

namespace gdpt {
int poisson_device(gdpt_ctx *c, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, double alpha, double *d_out);
int merge_device(gdpt_ctx *ctx, int w, int h, int ya, int yb, const double *d_cx0, const double *d_cx1, const double *d_cy0, const double *d_cy1, double *d_cx, double *d_cy);
}  // namespace gdpt

using namespace gdpt;

namespace {

constexpr int ROWS_WARPS = 4;          // rows per block in the row layout

// One warp per film row.  The per-sample arithmetic is sg::sample_pixel's, split over four lanes: what lane k adds to its
// own running sum is exactly the k-th of the four `rd.. = rd.. + ...` lines (:512-515), so the sums are the reference's.
__global__ void __launch_bounds__(32 * ROWS_WARPS) k_smallgdpt_rows(sg::Scene sc, int samps, double *c, double *cx0, double *cy0, double *cx1, double *cy1) {
    __shared__ sg::Path base[ROWS_WARPS];
    __shared__ double s_contrib[ROWS_WARPS][3], s_prob[ROWS_WARPS];
    __shared__ int s_ok[ROWS_WARPS];
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int y = blockIdx.x * ROWS_WARPS + wid;
    if (y >= sc.height || lane >= 4) return;                 // lanes 4..31 have no part in this layout
    constexpr unsigned MASK = 0xFu;
    sg::Path &path = base[wid];
    sg::Path off;
    uint64_t rng = sg::row_seed(y);
    const int dxs[4] = {-1, 0, 1, 0}, dys[4] = {0, 1, 0, -1};             // X0, Y0, X1, Y1 (:490-510)
    double *const dst[4] = {cx0, cy0, cx1, cy1};
    for (int x = 0; x < sc.width; x++) {
        sg::V r = sg::mk(0, 0, 0), rd = sg::mk(0, 0, 0);
        for (int s = 0; s < samps; s++) {
            if (lane == 0) {
                int ok = sg::generate_path(sc, x, y, rng, path) ? 1 : 0;
                if (ok) {
                    sg::V contrib; double prob;
                    sg::path_contrib_prob(sc, path, path.rnds[0], path.rnds[1], contrib, prob);
                    ok = prob > 0.0;
                    s_contrib[wid][0] = contrib.x; s_contrib[wid][1] = contrib.y; s_contrib[wid][2] = contrib.z; s_prob[wid] = prob;
                }
                s_ok[wid] = ok;
            }
            __syncwarp(MASK);
            if (s_ok[wid]) {
                const sg::V contrib = sg::mk(s_contrib[wid][0], s_contrib[wid][1], s_contrib[wid][2]);
                const double prob = s_prob[wid];
                if (lane == 0) r = r + (contrib * (1.0 / prob)) * (1.0 / (double)samps);
                sg::V cS = sg::mk(0, 0, 0); double wS = 1, jacobian;
                if (sg::shift_path(sc, x + dxs[lane], y + dys[lane], path, off, jacobian)) {
                    sg::V cc; double p;
                    sg::path_contrib_prob(sc, off, 0.0, 0.0, cc, p);
                    cS = cc * jacobian;
                    wS = prob / (prob + p * jacobian);
                }
                rd = rd + ((lane < 2) ? (contrib - cS) : (cS - contrib)) * (wS / (prob * (double)samps));
            }
            __syncwarp(MASK);                                 // the base path is rewritten by the next sample
        }
        const size_t i = ((size_t)(sc.height - y - 1) * sc.width + x) * 3;
        if (lane == 0) { c[i] = r.x; c[i + 1] = r.y; c[i + 2] = r.z; }
        double *d = dst[lane];
        d[i] = rd.x; d[i + 1] = rd.y; d[i + 2] = rd.z;
    }
}

__global__ void __launch_bounds__(128) k_smallgdpt_pixels(sg::Scene sc, int samps, int tiles_x, long long n_threads,
                                                          double *c, double *cx0, double *cy0, double *cx1, double *cy1) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_threads) return;
    const long long tile = t >> 5; const int lane = (int)(t & 31);
    const int x = (int)(tile % tiles_x) * 8 + (lane & 7), y = (int)(tile / tiles_x) * 4 + (lane >> 3);
    if (x >= sc.width || y >= sc.height) return;
    sg::Path path, off;
    uint64_t rng = sg::pixel_seed(x, y, sc.width);
    sg::render_pixel(sc, x, y, samps, rng, path, off, c, cx0, cy0, cx1, cy1);
}

}  // namespace

extern "C" int gdpt_smallgdpt_render(int device, int width, int height, int spp, int rng_mode, gdpt_images *out, double *render_ms) {
    if (render_ms) *render_ms = 0.0;
    if (width < 2 || height < 2 || width > 16384 || height > 16384) return fail(GDPT_ERR_ARG, "gdpt_smallgdpt_render: film %dx%d outside 2..16384", width, height);
    if (spp < 1 || spp > (1 << 20)) return fail(GDPT_ERR_ARG, "gdpt_smallgdpt_render: spp %d outside 1..2^20", spp);
    if (rng_mode != GDPT_SMALLGDPT_ROWS && rng_mode != GDPT_SMALLGDPT_PIXELS) return fail(GDPT_ERR_ARG, "gdpt_smallgdpt_render: unknown rng_mode %d", rng_mode);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(GDPT_ERR_CUDA, "no CUDA device: libgdpt has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(GDPT_ERR_ARG, "device out of range");
    GD_CUDA(cudaSetDevice(device));
    gdpt_ctx *c = new gdpt_ctx();                            // a bare context (no lajolla scene): stream, image buffers, Poisson plans
    c->device = device; memset(&c->stats, 0, sizeof(c->stats));
    for (auto &e : c->ev) e = nullptr;
    const int w = width, h = height;
    const size_t bytes = (size_t)w * h * 3 * sizeof(double);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    auto run = [&]() -> int {
        GD_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        c->own_stream = true;
        DevBuf *bufs[] = {&c->img, &c->cx0, &c->cy0, &c->cx1, &c->cy1, &c->cx, &c->cy, &c->out};
        for (auto *b : bufs) if (int r = b->ensure(bytes)) return r;
        GD_CUDA(cudaEventCreate(&e0)); GD_CUDA(cudaEventCreate(&e1));
        const sg::Scene sc = sg::make_scene(w, h);
        double *img = c->img.as<double>(), *x0 = c->cx0.as<double>(), *y0 = c->cy0.as<double>(), *x1 = c->cx1.as<double>(), *y1 = c->cy1.as<double>();
        GD_CUDA(cudaEventRecord(e0, c->stream));
        if (rng_mode == GDPT_SMALLGDPT_ROWS) {
            k_smallgdpt_rows<<<(h + ROWS_WARPS - 1) / ROWS_WARPS, 32 * ROWS_WARPS, 0, c->stream>>>(sc, spp, img, x0, y0, x1, y1);
        } else {
            const int tiles_x = (w + 7) / 8, tiles_y = (h + 3) / 4;
            const long long n = (long long)tiles_x * tiles_y * 32;
            k_smallgdpt_pixels<<<(unsigned)((n + 127) / 128), 128, 0, c->stream>>>(sc, spp, tiles_x, n, img, x0, y0, x1, y1);
        }
        GD_CUDA(cudaGetLastError());
        GD_CUDA(cudaEventRecord(e1, c->stream));
        if (int r = merge_device(c, w, h, 0, h, x0, x1, y0, y1, c->cx.as<double>(), c->cy.as<double>())) return r;       // :530-539
        if (int r = poisson_device(c, w, h, img, c->cx.as<double>(), c->cy.as<double>(), 0.04, c->out.as<double>())) return r;   // :541
        if (out) {
            if (out->primal) GD_CUDA(cudaMemcpyAsync(out->primal, img, bytes, cudaMemcpyDeviceToHost, c->stream));
            if (out->dx) GD_CUDA(cudaMemcpyAsync(out->dx, c->cx.p, bytes, cudaMemcpyDeviceToHost, c->stream));
            if (out->dy) GD_CUDA(cudaMemcpyAsync(out->dy, c->cy.p, bytes, cudaMemcpyDeviceToHost, c->stream));
            if (out->poisson) GD_CUDA(cudaMemcpyAsync(out->poisson, c->out.p, bytes, cudaMemcpyDeviceToHost, c->stream));
        }
        GD_CUDA(cudaStreamSynchronize(c->stream));
        float ms = 0.f;
        GD_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        if (render_ms) *render_ms = ms;
        return 0;
    };
    const int rc = run();
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    destroy_context(c);
    return rc;
}
