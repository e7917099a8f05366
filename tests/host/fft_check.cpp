// CPU check of the device functions of csrc/gdpt_poisson_fft.cuh (radix-4 forward / inverse passes, Bluestein DCT-I line
// transform): the header is compiled for the host with one emulated thread, the plan tables are built like the setup
// kernels build them, and the result is compared with the O(N^2) definition of REDFT00.  Prints the max relative error
// per length as JSON.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
struct double2 { double x, y; };
static inline double2 make_double2(double x, double y) { return double2{x, y}; }
struct Dim { int x; };
static Dim threadIdx{0}, blockDim{1};
#define __device__
#define __forceinline__ inline
#define __restrict__
#define __host__
static inline void __syncthreads() {}
template <typename T> static inline T __ldg(const T *p) { return *p; }
static inline double cospi(double x) { return cos(M_PI * x); }
static inline double sinpi(double x) { return sin(M_PI * x); }
#define GDPT_FFT_HOST_TEST
#include "gdpt_poisson_fft.cuh"
using namespace gdpt;

int main(int argc, char **argv) {
    printf("{");
    bool first = true;
    for (int a = 1; a < argc; a++) {
        const int N = atoi(argv[a]), M = N - 1;
        int P = 16, logP = 4; while (P < 2 * M - 1) { P <<= 1; logP++; }
        const int T = fft_twiddle_count(P, logP);
        std::vector<double2> tw(T), chirp(M), wpost(M), bhat(P), s(P);
        for (int i = 0; i < T; i++) {
            int off = 0, n = P;
            while (i >= off + (n >> 2)) { off += n >> 2; n >>= 2; }
            double t = 2.0 * (i - off) / n; tw[i] = make_double2(cospi(t), -sinpi(t));
        }
        for (int i = 0; i < M; i++) {
            long long q = ((long long)i * i) % (2LL * M);
            chirp[i] = make_double2(cospi((double)q / M), sinpi((double)q / M));
            wpost[i] = make_double2(cospi((double)i / M), -sinpi((double)i / M));
        }
        for (int i = 0; i < P; i++) { double2 v = make_double2(0, 0); if (i < M) v = chirp[i]; else if (P - i < M) v = chirp[P - i]; s[fft_sw(i)] = v; }   // the line buffer's layout: fft_sw
        fft_forward(s.data(), P, logP, tw.data());
        for (int i = 0; i < P; i++) bhat[i] = make_double2(s[fft_sw(i)].x / P, s[fft_sw(i)].y / P);
        // round trip of the two FFT routines
        std::vector<double2> r(P);
        srand(N);
        for (int i = 0; i < P; i++) r[i] = make_double2(rand() / (double)RAND_MAX - 0.5, rand() / (double)RAND_MAX - 0.5);
        std::vector<double2> r0 = r;
        fft_forward(r.data(), P, logP, tw.data()); fft_inverse(r.data(), P, logP, tw.data());
        double rt = 0; for (int i = 0; i < P; i++) rt = fmax(rt, fmax(fabs(r[i].x / P - r0[i].x), fabs(r[i].y / P - r0[i].y)));
        DctPlan pl{N, M, P, logP, tw.data(), chirp.data(), bhat.data(), wpost.data()};
        std::vector<double> x(N + 1), want(N);
        for (int n = 0; n < N; n++) x[n] = rand() / (double)RAND_MAX - 0.3;
        for (int k = 0; k < N; k++) {
            long double acc = x[0] + ((k & 1) ? -x[M] : x[M]);
            for (int n = 1; n < M; n++) acc += 2.0L * x[n] * cosl(M_PIl * (long double)(((long long)n * k) % (2LL * M)) / M);
            want[k] = (double)acc;
        }
        dct1_line(pl, x.data(), s.data(), tw.data());
        double err = 0, scale = 0;
        for (int k = 0; k < N; k++) { err = fmax(err, fabs(x[k] - want[k])); scale = fmax(scale, fabs(want[k])); }
        printf("%s\"%d\": {\"P\": %d, \"roundtrip\": %.3e, \"dct_rel_err\": %.3e}", first ? "" : ", ", N, P, rt, err / scale);
        first = false;
    }
    printf("}\n");
    return 0;
}
