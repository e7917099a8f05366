/* TEST INFRASTRUCTURE — CPU oracle, not part of the product.
 *
 * Unnormalised DCT-I (FFTW "REDFT00"), the transform the reference's
 * fourierSolve runs through fftw_plan_r2r_2d(h, w, buf, buf, FFTW_REDFT00,
 * FFTW_REDFT00, FFTW_ESTIMATE)  (/root/reference/src/render.cpp:190-192,227,242).
 * FFTW 3.3.10 is a third-party dependency vendored in the reference tree as
 * source (src/fftw-3.3.10, darwin-configured, needs its own cmake/autotools —
 * treated as unbuildable here); its published definition of REDFT00 is
 *
 *     X_k = x_0 + (-1)^k x_{N-1} + 2 * sum_{n=1}^{N-2} x_n cos(pi n k/(N-1))
 *
 * which equals the length-M DFT (M = 2(N-1)) of the even extension of x.
 * Two algorithms are provided so they can pin each other:
 *   gdpt_oracle_dct1_direct  — the O(N^2) definition, exact argument reduction
 *   gdpt_oracle_dct1_2d      — Bluestein chirp-z on a power-of-two FFT (any N),
 *                              two real rows packed per complex transform.
 * Both are cross-checked against scipy.fft.dctn(type=1) in tests/.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

void gdpt_oracle_dct1_direct(const double *x, double *X, int n) {
    int M = 2 * (n - 1);
    double *ct = (double *)malloc(sizeof(double) * (size_t)M);
    for (int j = 0; j < M; j++) ct[j] = cos(M_PI * (double)j / (double)(n - 1));
    for (int k = 0; k < n; k++) {
        double s = x[0] + ((k & 1) ? -x[n - 1] : x[n - 1]);
        for (int j = 1; j < n - 1; j++) s += 2.0 * x[j] * ct[(int)(((long long)j * k) % M)];
        X[k] = s;
    }
    free(ct);
}

/* iterative radix-2 complex FFT, sign = -1 forward, +1 inverse (unscaled) */
static void fft_pow2(double *re, double *im, int L, const double *wr, const double *wi, int sign) {
    for (int i = 1, j = 0; i < L; i++) {
        int bit = L >> 1;
        for (; j & bit; bit >>= 1) j ^= bit;
        j ^= bit;
        if (i < j) { double t = re[i]; re[i] = re[j]; re[j] = t; t = im[i]; im[i] = im[j]; im[j] = t; }
    }
    for (int len = 2; len <= L; len <<= 1) {
        int half = len >> 1, step = L / len;
        for (int i = 0; i < L; i += len)
            for (int j = 0; j < half; j++) {
                double c = wr[j * step], s = sign * wi[j * step];
                double ur = re[i + j], ui = im[i + j];
                double vr = re[i + j + half] * c - im[i + j + half] * s;
                double vi = re[i + j + half] * s + im[i + j + half] * c;
                re[i + j] = ur + vr; im[i + j] = ui + vi;
                re[i + j + half] = ur - vr; im[i + j + half] = ui - vi;
            }
    }
}

typedef struct {
    int n, M, L;
    double *wr, *wi;     /* twiddles exp(-2 pi i j / L), j < L/2 */
    double *cr, *ci;     /* chirp c[j] = exp(-i pi j^2 / M), j < M */
    double *Br, *Bi;     /* FFT of conj-chirp kernel, length L */
    double *ar, *ai;     /* work */
} dct1_plan;

static dct1_plan *plan_new(int n) {
    dct1_plan *p = (dct1_plan *)calloc(1, sizeof(dct1_plan));
    p->n = n; p->M = 2 * (n - 1);
    int L = 1; while (L < 2 * p->M - 1) L <<= 1;
    p->L = L;
    p->wr = (double *)malloc(sizeof(double) * L / 2 + 8); p->wi = (double *)malloc(sizeof(double) * L / 2 + 8);
    for (int j = 0; j < L / 2; j++) { p->wr[j] = cos(2 * M_PI * j / L); p->wi[j] = sin(2 * M_PI * j / L); }
    p->cr = (double *)malloc(sizeof(double) * p->M); p->ci = (double *)malloc(sizeof(double) * p->M);
    for (int j = 0; j < p->M; j++) {
        long long q = ((long long)j * j) % (2LL * p->M);   /* exact reduction of j^2 mod 2M */
        p->cr[j] = cos(M_PI * (double)q / p->M); p->ci[j] = -sin(M_PI * (double)q / p->M);
    }
    p->Br = (double *)calloc(L, sizeof(double)); p->Bi = (double *)calloc(L, sizeof(double));
    for (int j = 0; j < p->M; j++) {
        p->Br[j] = p->cr[j]; p->Bi[j] = -p->ci[j];
        if (j) { p->Br[L - j] = p->cr[j]; p->Bi[L - j] = -p->ci[j]; }
    }
    fft_pow2(p->Br, p->Bi, L, p->wr, p->wi, -1);
    p->ar = (double *)malloc(sizeof(double) * L); p->ai = (double *)malloc(sizeof(double) * L);
    return p;
}

static void plan_free(dct1_plan *p) {
    free(p->wr); free(p->wi); free(p->cr); free(p->ci); free(p->Br); free(p->Bi); free(p->ar); free(p->ai); free(p);
}

/* DCT-I of two real sequences x1,x2 (stride s) at once; x2 may be NULL. In place. */
static void dct1_pair(dct1_plan *p, double *x1, double *x2, int s) {
    int n = p->n, M = p->M, L = p->L;
    double *ar = p->ar, *ai = p->ai;
    memset(ar, 0, sizeof(double) * L); memset(ai, 0, sizeof(double) * L);
    for (int j = 0; j < M; j++) {
        int src = j < n ? j : M - j;
        double yr = x1[(size_t)src * s], yi = x2 ? x2[(size_t)src * s] : 0.0;
        ar[j] = yr * p->cr[j] - yi * p->ci[j];
        ai[j] = yr * p->ci[j] + yi * p->cr[j];
    }
    fft_pow2(ar, ai, L, p->wr, p->wi, -1);
    for (int j = 0; j < L; j++) {
        double r = ar[j] * p->Br[j] - ai[j] * p->Bi[j];
        double i = ar[j] * p->Bi[j] + ai[j] * p->Br[j];
        ar[j] = r; ai[j] = i;
    }
    fft_pow2(ar, ai, L, p->wr, p->wi, +1);
    double inv = 1.0 / L;
    for (int k = 0; k < n; k++) {
        double r = ar[k] * inv, i = ai[k] * inv;
        x1[(size_t)k * s] = r * p->cr[k] - i * p->ci[k];
        if (x2) x2[(size_t)k * s] = r * p->ci[k] + i * p->cr[k];
    }
}

/* In-place 2-D DCT-I of a row-major h x w array (REDFT00 on both axes). */
void gdpt_oracle_dct1_2d(double *buf, int h, int w) {
    dct1_plan *pw = plan_new(w);
    for (int y = 0; y < h; y += 2)
        dct1_pair(pw, buf + (size_t)y * w, y + 1 < h ? buf + (size_t)(y + 1) * w : NULL, 1);
    plan_free(pw);
    dct1_plan *ph = plan_new(h);
    for (int x = 0; x < w; x += 2)
        dct1_pair(ph, buf + x, x + 1 < w ? buf + x + 1 : NULL, w);
    plan_free(ph);
}

void gdpt_oracle_dct1_1d(double *x, int n) {
    dct1_plan *p = plan_new(n);
    dct1_pair(p, x, NULL, 1);
    plan_free(p);
}
