// DCT-I (FFTW's REDFT00, what fourierSolve executes twice per channel: /root/reference/src/render.cpp:190-192,227,242)
// in O(N log N): the replacement for the dense-matrix form of gdpt_poisson.cuh on films of 16 pixels a side and more.
//
//   X_k = x_0 + (-1)^k x_M + 2 sum_{n=1}^{M-1} x_n cos(pi n k / M),   M = N - 1,  k = 0..M
//
// is the DFT of length 2M of the even extension y of x.  Because y is real, that DFT follows from ONE complex DFT of
// length M of z_n = y_2n + i y_2n+1 (the usual even/odd split, stable in floating point).  2M has large prime factors at
// every film size of interest (1919*2, 1079*2, 3839*2, 2159*2: SURVEY 7.2-4), so the length-M DFT is evaluated with
// Bluestein's chirp-z identity  Z_k = conj(w_k) * sum_n (z_n conj(w_n)) w_(k-n),  w_n = exp(i pi n^2 / M),  as a circular
// convolution of power-of-two length P >= 2M - 1:  IFFT_P( FFT_P(a) * FFT_P(b) ).
//
// One thread block owns one line (a row, or a column read with its stride): the line is staged in shared memory, both
// power-of-two FFTs run there as radix-4 passes (decimation in frequency forward, the exact structural inverse —
// decimation in time — backward, so no digit-reversal pass is ever needed: the filter's transform is produced by the same
// forward routine and therefore in the same order), and the column pass keeps the line on chip across "forward DCT,
// spectral divide (render.cpp:231-239), inverse DCT".  FP64 throughout; the chirp angles are reduced exactly in integers
// (n^2 mod 2M) before cospi / sinpi.
#pragma once
#ifndef GDPT_FFT_HOST_TEST      // tests/host/fft_check.cpp compiles the device functions below for the host (one "thread")
#include <cuda_runtime.h>
#endif
#include <math.h>

namespace gdpt {

struct DctPlan {
    int N, M, P, logP;
    const double2 *tw;      // [T]   per radix-4 pass, contiguous: exp(-2 pi i j / n) for j < n/4, n = P, P/4, ... (T = fft_twiddle_count(P))
    const double2 *chirp;   // [M]   w_n = exp(i pi n^2 / M)
    const double2 *bhat;    // [P]   FFT_P of the circular chirp filter, in the forward routine's output order, times 1/P
    const double2 *wpost;   // [M]   exp(-i pi k / M)
};

// twiddles of all radix-4 passes of a length-P transform, pass after pass (forward order): sum of n/4 over n = P, P/4, ...
__host__ __device__ inline int fft_twiddle_count(int P, int logP) { int t = 0; for (int s = 0, n = P; s < logP / 2; s++, n >>= 2) t += n >> 2; return t; }

constexpr int FFT_THREADS = 512;      // 16 warps per block, two blocks per SM at 1080p (80 KB each): the passes are latency / barrier bound

__device__ __forceinline__ double2 cmul(double2 a, double2 b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ double2 cmulc(double2 a, double2 b) { return make_double2(a.x * b.x + a.y * b.y, a.y * b.x - a.x * b.y); }   // a * conj(b)
__device__ __forceinline__ double2 cadd(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ double2 csub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }

// Where element e of the line lives in shared memory.  An element is 16 bytes, so a warp's access is served a quarter-warp
// (8 lanes, 8 sixteen-byte banks) at a time.  In the natural layout the passes with butterfly stride 4, 2 and 1 (and the final
// radix-2 pass of an odd log2 P) put 2 to 4 lanes of a quarter-warp on one bank: 34 % of the kernel's shared-memory wavefronts
// were replays (ncu, l1tex__data_bank_conflicts_pipe_lsu_mem_shared, profiles/r2_k_dct_cols_final.txt).  XOR-ing the low three
// index bits with (x ^ x<<1), x = bits 3..4, is the cheapest linear map that makes EVERY access pattern of both FFT directions
// at P = 2048 / 4096 / 8192 and the contiguous sweeps conflict-free (exhaustive search over the 3-bit XOR masks of bits 3..6).
__host__ __device__ __forceinline__ int fft_sw(int e) { const int x = (e >> 3) & 3; return e ^ (x ^ (x << 1)); }

// In-place forward FFT of s[0..P) (natural order in, digit-reversed out), all threads of the block.
// tw: the per-pass twiddle table (DctPlan::tw), staged in shared memory by the kernels: only w^j is read, w^2j and w^3j are
// formed in registers — the FP64 pipe idles while the shared-memory / L1 pipe is the busy one (ncu: l1tex 94 %).
// first_stage = 1: the first radix-4 pass has already been done (dct1_line forms it straight from the real line).
__device__ inline void fft_forward(double2 *s, int P, int logP, const double2 *__restrict__ tw, int first_stage = 0) {
    int n = P, ln = logP, off = 0;        // block length of the pass, its log2, start of the pass's twiddles
    if (first_stage) { off = n >> 2; n >>= 2; ln -= 2; }
    for (int stage = first_stage; stage < logP / 2; stage++, off += n >> 2, n >>= 2, ln -= 2) {
        const int q4 = n >> 2, lq = ln - 2;
        for (int idx = threadIdx.x; idx < (P >> 2); idx += blockDim.x) {
            const int blk = idx >> lq, j = idx & (q4 - 1), base = (blk << ln) + j;
            const double2 a = s[fft_sw(base)], b = s[fft_sw(base + q4)], c = s[fft_sw(base + 2 * q4)], d = s[fft_sw(base + 3 * q4)];
            const double2 t0 = cadd(a, c), t1 = csub(a, c), t2 = cadd(b, d), t3 = csub(b, d);
            const double2 y1 = make_double2(t1.x + t3.y, t1.y - t3.x), y3 = make_double2(t1.x - t3.y, t1.y + t3.x);
            s[fft_sw(base)] = cadd(t0, t2);
            if (j == 0) { s[fft_sw(base + q4)] = y1; s[fft_sw(base + 2 * q4)] = csub(t0, t2); s[fft_sw(base + 3 * q4)] = y3; }
            else {
                const double2 w1 = tw[off + j], w2 = cmul(w1, w1), w3 = cmul(w2, w1);
                s[fft_sw(base + q4)] = cmul(y1, w1);
                s[fft_sw(base + 2 * q4)] = cmul(csub(t0, t2), w2);
                s[fft_sw(base + 3 * q4)] = cmul(y3, w3);
            }
        }
        __syncthreads();
    }
    if (logP & 1) {
        for (int idx = threadIdx.x; idx < (P >> 1); idx += blockDim.x) {
            const double2 a = s[fft_sw(2 * idx)], b = s[fft_sw(2 * idx + 1)];
            s[fft_sw(2 * idx)] = cadd(a, b); s[fft_sw(2 * idx + 1)] = csub(a, b);
        }
        __syncthreads();
    }
}

// The structural inverse of fft_forward (digit-reversed in, natural out), unscaled: fft_inverse(fft_forward(x)) = P x.
// keep: only outputs with index < keep are needed (the last pass does not store the others);
// pre (may be null): element e is multiplied by pre[e] as the first stage loads it — the filter product of the circular
// convolution rides on the first pass instead of being a shared-memory pass (and a barrier) of its own.
__device__ inline void fft_inverse(double2 *s, int P, int logP, const double2 *__restrict__ tw, const double2 *__restrict__ pre = nullptr, int keep = 0x7fffffff) {
    if (logP & 1) {
        for (int idx = threadIdx.x; idx < (P >> 1); idx += blockDim.x) {
            double2 a = s[fft_sw(2 * idx)], b = s[fft_sw(2 * idx + 1)];
            if (pre) { a = cmul(a, __ldg(pre + 2 * idx)); b = cmul(b, __ldg(pre + 2 * idx + 1)); }
            s[fft_sw(2 * idx)] = cadd(a, b); s[fft_sw(2 * idx + 1)] = csub(a, b);
        }
        __syncthreads();
        pre = nullptr;
    }
    int n = (logP & 1) ? 8 : 4, ln = (logP & 1) ? 3 : 2, off = fft_twiddle_count(P, logP);
    for (int stage = 0; stage < logP / 2; stage++, n <<= 2, ln += 2) {
        const int q4 = n >> 2, lq = ln - 2;
        off -= q4;
        for (int idx = threadIdx.x; idx < (P >> 2); idx += blockDim.x) {
            const int blk = idx >> lq, j = idx & (q4 - 1), base = (blk << ln) + j;
            double2 u0 = s[fft_sw(base)], u1 = s[fft_sw(base + q4)], u2 = s[fft_sw(base + 2 * q4)], u3 = s[fft_sw(base + 3 * q4)];
            if (stage == 0 && pre) { u0 = cmul(u0, __ldg(pre + base)); u1 = cmul(u1, __ldg(pre + base + q4)); u2 = cmul(u2, __ldg(pre + base + 2 * q4)); u3 = cmul(u3, __ldg(pre + base + 3 * q4)); }
            if (j != 0) { const double2 w1 = tw[off + j], w2 = cmul(w1, w1), w3 = cmul(w2, w1); u1 = cmulc(u1, w1); u2 = cmulc(u2, w2); u3 = cmulc(u3, w3); }
            const double2 t0 = cadd(u0, u2), t1 = csub(u0, u2), t2 = cadd(u1, u3), t3 = csub(u1, u3);
            s[fft_sw(base)] = cadd(t0, t2);
            if (n < P || base + q4 < keep) s[fft_sw(base + q4)] = make_double2(t1.x - t3.y, t1.y + t3.x);      // u0 + i u1 - u2 - i u3
            if (n < P || base + 2 * q4 < keep) {
                s[fft_sw(base + 2 * q4)] = csub(t0, t2);
                s[fft_sw(base + 3 * q4)] = make_double2(t1.x + t3.y, t1.y - t3.x);  // u0 - i u1 - u2 + i u3
            }
        }
        __syncthreads();
    }
}

// xs[0..N) (real, shared memory) -> its DCT-I in place.  s: P complex of shared memory.
__device__ inline void dct1_line(const DctPlan &pl, double *xs, double2 *s, const double2 *tw) {
    const int M = pl.M, P = pl.P;
    // z_n = (y_2n + i y_2n+1) conj(w_n) for n < M, zero up to P — and since M <= P/2 the upper half of the padded line is all
    // zeros, so the first radix-4 pass (block length P: inputs j, j + P/4, j + P/2, j + 3P/4) has c = d = 0 and is formed here
    // from the real line directly: no store of the padded line, no pass over it
    auto z = [&](int n) {
        if (n >= M) return make_double2(0.0, 0.0);
        const int e = 2 * n, o = 2 * n + 1;
        return cmulc(make_double2(xs[e <= M ? e : 2 * M - e], xs[o <= M ? o : 2 * M - o]), __ldg(pl.chirp + n));
    };
    const int q4 = P >> 2;
    for (int j = threadIdx.x; j < q4; j += blockDim.x) {
        const double2 a = z(j), b = z(j + q4);
        const double2 y1 = make_double2(a.x + b.y, a.y - b.x), y3 = make_double2(a.x - b.y, a.y + b.x), y2 = csub(a, b);
        s[fft_sw(j)] = cadd(a, b);
        if (j == 0) { s[fft_sw(q4)] = y1; s[fft_sw(2 * q4)] = y2; s[fft_sw(3 * q4)] = y3; }
        else {
            const double2 w1 = tw[j], w2 = cmul(w1, w1), w3 = cmul(w2, w1);
            s[fft_sw(j + q4)] = cmul(y1, w1);
            s[fft_sw(j + 2 * q4)] = cmul(y2, w2);
            s[fft_sw(j + 3 * q4)] = cmul(y3, w3);
        }
    }
    __syncthreads();
    fft_forward(s, P, pl.logP, tw, 1);
    fft_inverse(s, P, pl.logP, tw, pl.bhat, M);
    for (int k = threadIdx.x; k < M; k += blockDim.x) {
        const int km = k == 0 ? 0 : M - k;                                        // Z_k = s_k conj(w_k), formed where it is used
        const double2 zk = cmulc(s[fft_sw(k)], __ldg(pl.chirp + k)), zm = cmulc(s[fft_sw(km)], __ldg(pl.chirp + km)), wp = __ldg(pl.wpost + k);
        const double ex = 0.5 * (zk.x + zm.x), ey = 0.5 * (zk.y - zm.y);          // E = (Z_k + conj Z_(M-k)) / 2
        const double ox = 0.5 * (zk.y + zm.y), oy = -0.5 * (zk.x - zm.x);         // O = (Z_k - conj Z_(M-k)) / (2i)
        (void)ey;
        xs[k] = ex + wp.x * ox - wp.y * oy;
        if (k == 0) xs[M] = zk.x - zk.y;                                          // Y_M = E_0 - O_0
    }
    __syncthreads();
}

#ifndef GDPT_FFT_HOST_TEST
enum { DCT_STORE = 0, DCT_COLUMNS = 1, DCT_SCALE_INTERLEAVE = 2 };
struct DctArgs {
    int mode;
    int count;              // lines
    int per_plane;          // lines per channel plane (rows: H, columns: W)
    long long plane_stride; // elements between channel planes
    long long line_stride;  // elements between consecutive lines of a plane (rows: W, columns: 1)
    long long elem_stride;  // elements between consecutive samples of a line (rows: 1, columns: W)
    // DCT_COLUMNS: spectral divide between the two transforms (render.cpp:229-239)
    int img_w, img_h; double alpha; const double *dc;
    // DCT_SCALE_INTERLEAVE: out[(y * img_w + x) * 3 + ch] = value / scale (render.cpp:243-247)
    double scale; double *out;
};

static __global__ void __launch_bounds__(FFT_THREADS, 2) k_dct1_fft(DctPlan pl, DctArgs a, const double *in, double *dst) {   // in == dst is allowed (a block reads its whole line before it writes it)
    extern __shared__ double2 smem_fft[];
    double2 *s = smem_fft;
    const int N = pl.N, T = fft_twiddle_count(pl.P, pl.logP);
    double2 *twsm = smem_fft + pl.P;                                       // the pass twiddles, staged once per block
    double *xs = reinterpret_cast<double *>(smem_fft + pl.P + T);
    for (int i = threadIdx.x; i < T; i += blockDim.x) twsm[i] = pl.tw[i];
    __syncthreads();
    for (int line = blockIdx.x; line < a.count; line += gridDim.x) {
        const int ch = line / a.per_plane, l = line - ch * a.per_plane;
        const long long base = ch * a.plane_stride + l * a.line_stride;
        for (int n = threadIdx.x; n < N; n += blockDim.x) xs[n] = in[base + n * a.elem_stride];
        __syncthreads();
        dct1_line(pl, xs, s, twsm);
        if (a.mode == DCT_COLUMNS) {
            // l = image column x, sample index = frequency row y: F = H / (alpha - (float)(lapY + lapX)); F(0,0) = DC of the primal
            const double lapx = 2.0 * cos(3.14159265358979323846 * l / (a.img_w - 1));
            for (int y = threadIdx.x; y < N; y += blockDim.x) {
                const double lapy = -4.0 + (2.0 * cos(3.14159265358979323846 * y / (a.img_h - 1)));
                const float resp = (float)(lapy + lapx);
                double v = xs[y] / (a.alpha - resp);
                if (y == 0 && l == 0) v = a.dc[ch];
                xs[y] = v;
            }
            __syncthreads();
            dct1_line(pl, xs, s, twsm);
        }
        if (a.mode == DCT_SCALE_INTERLEAVE) {
            for (int n = threadIdx.x; n < N; n += blockDim.x) a.out[((long long)l * a.img_w + n) * 3 + ch] = xs[n] / a.scale;
        } else {
            for (int n = threadIdx.x; n < N; n += blockDim.x) dst[base + n * a.elem_stride] = xs[n];
        }
        __syncthreads();
    }
}

// Plan tables for one line length.  tw, chirp, wpost first (any grid), then the filter transform by ONE block.
static __global__ void k_dct_plan_tables(int M, int P, int logP, double2 *tw, double2 *chirp, double2 *wpost) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < fft_twiddle_count(P, logP)) {        // entry i belongs to the pass with block length n: exp(-2 pi i j / n)
        int off = 0, n = P;
        while (i >= off + (n >> 2)) { off += n >> 2; n >>= 2; }
        const double t = 2.0 * (double)(i - off) / (double)n;
        tw[i] = make_double2(cospi(t), -sinpi(t));
    }
    if (i < M) {
        const long long q = ((long long)i * i) % (2LL * M);
        const double t = (double)q / (double)M;
        chirp[i] = make_double2(cospi(t), sinpi(t));
        const double u = (double)i / (double)M;
        wpost[i] = make_double2(cospi(u), -sinpi(u));
    }
}
static __global__ void __launch_bounds__(FFT_THREADS) k_dct_plan_filter(int M, int P, int logP, const double2 *tw, const double2 *chirp, double2 *bhat) {
    extern __shared__ double2 smem_fft[];
    double2 *s = smem_fft;
    for (int i = threadIdx.x; i < P; i += blockDim.x) {        // (the twiddles are read from global memory here: once per size)
        double2 v = make_double2(0.0, 0.0);
        if (i < M) v = chirp[i];
        else if (P - i < M) v = chirp[P - i];       // b_(-m) = w_m
        s[fft_sw(i)] = v;
    }
    __syncthreads();
    fft_forward(s, P, logP, tw);
    const double inv = 1.0 / (double)P;
    for (int i = threadIdx.x; i < P; i += blockDim.x) { const double2 v = s[fft_sw(i)]; bhat[i] = make_double2(v.x * inv, v.y * inv); }
}

#endif  // GDPT_FFT_HOST_TEST

} // namespace gdpt
