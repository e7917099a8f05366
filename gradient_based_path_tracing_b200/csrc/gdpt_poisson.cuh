// Screened-Poisson reconstruction on the GPU: the replacement for fourierSolve
// (/root/reference/src/render.cpp:172-254) and for FFTW's REDFT00 plan it executes.
//
//   k_merge          cx = cx0 + cx1(x-1), cy = cy0 + cy1(y-1)            render.cpp:340-350
//   k_poisson_rhs    rhs = a*u - Dx - Dy, per-channel DC partial sums     render.cpp:202-225
//   k_dc_reduce      fixed-order final reduction of the DC sums (deterministic)
//   k_dct_matrix     D[n][k] = w_n cos(pi n k/(N-1)), w = 1 at the ends, 2 inside; argument
//                    reduced exactly in integers (n*k mod 2(N-1)) before cospi
//   k_dgemm          FP64 tiled GEMM (64x64x16 tiles staged in shared memory, 4x4 register
//                    blocks) that applies the 1-D DCT-I along rows / columns; epilogues fuse
//                    the spectral divide + DC replacement (render.cpp:231-239) and the final
//                    1/(4(W-1)(H-1)) scale + RGB re-interleave (render.cpp:243-247).
// This is the direct O(N^2)-per-line form (exact for every N: 2(N-1) has large prime
// factors at all the film sizes of interest); it is FP64-FMA bound.
// Work layout: planar [3][H][W] doubles.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

namespace gdpt {

// rows [ya, yb) of the full-size images (the single-GPU solve passes 0, h)
static __global__ void k_merge(int w, int h, int ya, int yb, const double *cx0, const double *cx1, const double *cy0, const double *cy1, double *cx, double *cy) {
    (void)h;
    size_t n = (size_t)w * yb * 3;
    for (size_t i = (size_t)w * ya * 3 + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        size_t pix = i / 3;
        int x = (int)(pix % w), y = (int)(pix / w);
        cx[i] = x == 0 ? cx0[i] : cx0[i] + cx1[i - 3];
        cy[i] = y == 0 ? cy0[i] : cy0[i] + cy1[i - (size_t)w * 3];
    }
}

constexpr int RHS_BLOCK = 256;
// u, gx, gy are full-size interleaved images; rows [y0, y1) are turned into the planar band buffer
// buf[3][y1-y0][w] (the single-GPU solve passes 0, h; a rank of the distributed solve its own band, whose
// gy must be valid up to row y1 when y1 < h)
static __global__ void __launch_bounds__(RHS_BLOCK) k_poisson_rhs(int w, int h, int y0, int y1, const double *u, const double *gx, const double *gy, double alpha,
                                                           double *buf, double *partial /* [gridDim.x][3] */) {
    __shared__ double red[3][RHS_BLOCK];
    double dc[3] = {0, 0, 0};
    size_t npix = (size_t)w * (y1 - y0);
    for (size_t q = blockIdx.x * (size_t)RHS_BLOCK + threadIdx.x; q < npix; q += (size_t)gridDim.x * RHS_BLOCK) {
        size_t p = q + (size_t)w * y0;
        int x = (int)(p % w), y = (int)(p / w);
        bool ix = x > 0 && x < w - 1, iy = y > 0 && y < h - 1;
        double m = 1.0;
        if (ix) m *= 2.0;
        if (iy) m *= 2.0;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            size_t a = p * 3 + c;
            double uv = u[a];
            dc[c] += m * uv;
            double v = alpha * uv;
            if (ix) v -= (gx[a + 3] - gx[a]); else v -= (-2.0 * gx[a]);
            if (iy) v -= (gy[a + (size_t)w * 3] - gy[a]); else v -= (-2.0 * gy[a]);
            buf[(size_t)c * npix + q] = v;
        }
    }
    for (int c = 0; c < 3; c++) red[c][threadIdx.x] = dc[c];
    __syncthreads();
    for (int s = RHS_BLOCK / 2; s > 0; s >>= 1) {
        if (threadIdx.x < s) for (int c = 0; c < 3; c++) red[c][threadIdx.x] += red[c][threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) for (int c = 0; c < 3; c++) partial[blockIdx.x * 3 + c] = red[c][0];
}

// one warp per channel (launch with >= 96 threads): lane l adds partials l, l+32, ... in that order, then a shuffle tree —
// a fixed order, so the DC sums are the same on every run
static __global__ void k_dc_reduce(int nblocks, const double *partial, double *dc) {
    const int c = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (c >= 3) return;
    double s = 0;
    for (int b = lane; b < nblocks; b += 32) s += partial[b * 3 + c];
    for (int d = 16; d > 0; d >>= 1) s += __shfl_down_sync(0xffffffffu, s, d);
    if (lane == 0) dc[c] = s;
}

// D[n][k] (transpose=0) or D[k][n] (transpose=1) = w_n cos(pi n k/(N-1))
static __global__ void k_dct_matrix(int N, int transpose, double *D) {
    size_t total = (size_t)N * N;
    long long M = 2LL * (N - 1);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        int r = (int)(i / N), c = (int)(i % N);
        int n = transpose ? c : r, k = transpose ? r : c;
        long long q = ((long long)n * k) % M;
        double wgt = (n == 0 || n == N - 1) ? 1.0 : 2.0;
        D[i] = wgt * cospi((double)q / (double)(N - 1));
    }
}

enum { EPI_STORE = 0, EPI_DIVIDE = 1, EPI_SCALE_INTERLEAVE = 2, EPI_PACK = 3 };
// Distributed solve: a rank's rows x W band of the three channels laid out chunk-major by COLUMN band, the
// chunk of column band s holding [rows][3][w_s] — i.e. exactly the message rank s receives in the all-to-all
// transpose, and (stacked over the sending ranks) a row-major [H][3*w_s] matrix on the receiving side.
// Column bands are band_w wide except the last.
struct Banded {
    int n_bands, band_w, rows, W;
    __host__ __device__ size_t addr(int ch, int y, int col) const {
        int s = col / band_w; if (s > n_bands - 1) s = n_bands - 1;
        int x0 = s * band_w, ws = (s == n_bands - 1) ? W - x0 : band_w;
        return (size_t)rows * 3 * x0 + (size_t)y * 3 * ws + (size_t)ch * ws + (col - x0);
    }
};
struct Epilogue {
    int mode;
    int img_w, img_h;   // full image size (spectral divide); img_h = rows of the band for the interleaving / packing epilogues
    double alpha, scale;
    const double *dc;   // [3]
    double *out;        // interleaved RGB for EPI_SCALE_INTERLEAVE
    int ch_w, x0;       // EPI_DIVIDE on a column band: columns are [3][ch_w] side by side starting at image column x0 (ch_w = 0: channel = blockIdx.z)
    Banded band;        // EPI_PACK store / BANDED_A load
};

constexpr int GT = 64, GK = 16;
// C[z] (MxN, ldc) = A (MxK, lda; shared by the batch when strideA == 0) * B[z] (KxN, ldb)
template <bool BANDED_A = false>
static __global__ void __launch_bounds__(256) k_dgemm(int M, int N, int K, const double *A, int lda, long long strideA,
                                               const double *B, int ldb, long long strideB, double *C, int ldc, long long strideC, Epilogue ep) {
    __shared__ double As[2][GK][GT + 1];
    __shared__ double Bs[2][GK][GT];
    const int z = blockIdx.z;
    A += z * strideA; B += z * strideB; C += z * strideC;
    const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
    const int row0 = blockIdx.y * GT, col0 = blockIdx.x * GT;
    double acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.0;
    // loaders: A tile 64 rows x 16 k -> 1024 elements, 4 per thread; B tile 16 k x 64 cols
    const int a_r = threadIdx.x / 4, a_k = (threadIdx.x % 4) * 4;
    const int b_k = threadIdx.x / 16, b_c = (threadIdx.x % 16) * 4;
    auto load = [&](int buf, int k0) {
#pragma unroll
        for (int q = 0; q < 4; q++) {
            int r = row0 + a_r, k = k0 + a_k + q;
            if (BANDED_A) As[buf][a_k + q][a_r] = (r < M && k < K) ? A[ep.band.addr(r / ep.band.rows, r % ep.band.rows, k)] : 0.0;
            else As[buf][a_k + q][a_r] = (r < M && k < K) ? A[(size_t)r * lda + k] : 0.0;
        }
#pragma unroll
        for (int q = 0; q < 4; q++) {
            int k = k0 + b_k, c = col0 + b_c + q;
            Bs[buf][b_k][b_c + q] = (k < K && c < N) ? B[(size_t)k * ldb + c] : 0.0;
        }
    };
    int nk = (K + GK - 1) / GK;
    load(0, 0);
    __syncthreads();
    for (int t = 0; t < nk; t++) {
        int cur = t & 1;
        if (t + 1 < nk) load(cur ^ 1, (t + 1) * GK);
#pragma unroll
        for (int k = 0; k < GK; k++) {
            double a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; i++) a[i] = As[cur][k][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; j++) b[j] = Bs[cur][k][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) acc[i][j] = fma(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int r = row0 + ty * 4 + i;
        if (r >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            int c = col0 + tx * 4 + j;
            if (c >= N) continue;
            double v = acc[i][j];
            if (ep.mode == EPI_STORE) {
                C[(size_t)r * ldc + c] = v;
            } else if (ep.mode == EPI_DIVIDE) {
                // r = y, c = x of channel z (whole image) or column ch*ch_w + (x - x0) of a column band
                int x = ep.ch_w ? ep.x0 + c % ep.ch_w : c, ch = ep.ch_w ? c / ep.ch_w : z;
                double lapx = 2.0 * cos(3.14159265358979323846 * x / (ep.img_w - 1));
                double lapy = -4.0 + (2.0 * cos(3.14159265358979323846 * r / (ep.img_h - 1)));
                float resp = (float)(lapy + lapx);
                v = v / (ep.alpha - resp);
                if (r == 0 && x == 0) v = ep.dc[ch];
                C[(size_t)r * ldc + c] = v;
            } else if (ep.mode == EPI_PACK) {
                C[ep.band.addr(r / ep.band.rows, r % ep.band.rows, c)] = v;
            } else {
                // rows of the planar [3][H] stack: r = ch*H + y
                int ch = r / ep.img_h, y = r % ep.img_h;
                ep.out[((size_t)y * ep.img_w + c) * 3 + ch] = v / ep.scale;
            }
        }
    }
}

} // namespace gdpt
