// L1 reconstruction of a gradient-domain render by iteratively reweighted least squares (Kettunen et al. 2015
// section 5.2, the paper handouts/final_project.tex:109-111 cites).  NO counterpart in the reference's code: both of its programs
// reconstruct with the L2 screened-Poisson solve only (fourierSolve, src/render.cpp:172-254; small_gdpt.cpp:355-436) — parity
// for this function is unpinned; its CPU statement is oracle/poisson_l1.py (exact sparse solves).
//
//   E(I) = alpha sum_p wU_p |I_p - u_p|^2 + sum_{x>=1} wX_p |I_p - I_(p-x) - gx_p|^2 + sum_{y>=1} wY_p |I_p - I_(p-y) - gy_p|^2
//
// gx(x, y) estimates I(x, y) - I(x-1, y), gy(x, y) estimates I(x, y) - I(x, y-1): the merged gradients of render.cpp:340-350;
// column 0 of gx and row 0 of gy are not used.  Weights: wX, wY = 1 / (eps + r), r = L1 norm over the channels of the term's
// current residual, one weight per pixel and term; eps = epsilon * mean|u|; wU = 1 unless params->reweight_primal (with a
// quadratic primal term the image keeps the primal's mean exactly — sum_p of D^T W (D I - g) telescopes to zero — where the
// all-L1 energy pulls every pixel towards its median, which darkens a path-traced primal whose samples are mostly zero).
// Each reweighting is followed by a Jacobi-preconditioned
// conjugate-gradient solve of the weighted normal equations, started from the current image.  Everything stays on the device:
// the CG scalars (step lengths, the convergence flag) live in device memory and the kernels read them, so one reconstruction is
// a fixed sequence of launches without a host read-back; the reductions are two-stage and fixed-order (bit-reproducible).
// Images are the product's interleaved [h][w][3] doubles; the stencils are HBM-bound streaming kernels (5 launches per CG step).
#include "gdpt_context.cuh"

namespace gdpt {
int poisson_device(gdpt_ctx *c, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, double alpha, double *d_out);
}
using namespace gdpt;

namespace {

constexpr int L1_BLOCK = 256;
constexpr int L1_GRID = 148 * 4;

struct L1Scalars {             // device
    double rz, pq, rr, bb, step, beta, mean_abs_u, eps;
    int done, iterations;
};

struct L1Geom { int w, h; long long n; double alpha; };

// fixed-order block reduction of up to two values; partial[blockIdx.x * 2 + k]
__device__ __forceinline__ void block_reduce2(double a, double b, double *partial) {
    __shared__ double sa[L1_BLOCK / 32], sb[L1_BLOCK / 32];
    for (int o = 16; o > 0; o >>= 1) { a += __shfl_down_sync(0xffffffffu, a, o); b += __shfl_down_sync(0xffffffffu, b, o); }
    if ((threadIdx.x & 31) == 0) { sa[threadIdx.x >> 5] = a; sb[threadIdx.x >> 5] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double ta = 0, tb = 0;
        for (int k = 0; k < L1_BLOCK / 32; k++) { ta += sa[k]; tb += sb[k]; }
        partial[blockIdx.x * 2] = ta; partial[blockIdx.x * 2 + 1] = tb;
    }
    __syncthreads();
}

// the weights of pixel p's five stencil arms and its diagonal entry
struct Arms { double l, r, u, d, diag; };
__device__ __forceinline__ Arms arms_of(const L1Geom &g, int x, int y, long long p, const double *wU, const double *wX, const double *wY) {
    Arms a;
    a.l = x >= 1 ? wX[p] : 0.0;
    a.r = x < g.w - 1 ? wX[p + 1] : 0.0;
    a.u = y >= 1 ? wY[p] : 0.0;
    a.d = y < g.h - 1 ? wY[p + g.w] : 0.0;
    a.diag = g.alpha * wU[p] + a.l + a.r + a.u + a.d;
    return a;
}
__device__ __forceinline__ double apply_at(const L1Geom &g, const Arms &a, long long e, int x, int y, const double *v) {
    double s = a.diag * v[e];
    if (x >= 1) s -= a.l * v[e - 3];
    if (x < g.w - 1) s -= a.r * v[e + 3];
    if (y >= 1) s -= a.u * v[e - 3LL * g.w];
    if (y < g.h - 1) s -= a.d * v[e + 3LL * g.w];
    return s;
}

__global__ void __launch_bounds__(L1_BLOCK) k_l1_mean_abs(L1Geom g, const double *u, double *partial) {
    double s = 0;
    for (long long e = blockIdx.x * (long long)L1_BLOCK + threadIdx.x; e < g.n; e += (long long)gridDim.x * L1_BLOCK) s += fabs(u[e]);
    block_reduce2(s, 0.0, partial);
}

// uniform = 1: all weights 1 (the L2 problem)
__global__ void __launch_bounds__(L1_BLOCK) k_l1_weights(L1Geom g, int uniform, int reweight_primal, const L1Scalars *sc, const double *I, const double *u, const double *gx, const double *gy,
                                                          double *wU, double *wX, double *wY) {
    const long long np = (long long)g.w * g.h;
    const double eps = sc->eps;
    for (long long p = blockIdx.x * (long long)L1_BLOCK + threadIdx.x; p < np; p += (long long)gridDim.x * L1_BLOCK) {
        if (uniform) { wU[p] = 1.0; wX[p] = 1.0; wY[p] = 1.0; continue; }
        const int x = (int)(p % g.w), y = (int)(p / g.w);
        const long long e = p * 3;
        double ru = 0, rx = 0, ry = 0;
        for (int c = 0; c < 3; c++) {
            ru += fabs(I[e + c] - u[e + c]);
            if (x >= 1) rx += fabs(I[e + c] - I[e + c - 3] - gx[e + c]);
            if (y >= 1) ry += fabs(I[e + c] - I[e + c - 3LL * g.w] - gy[e + c]);
        }
        wU[p] = reweight_primal ? 1.0 / (eps + ru) : 1.0; wX[p] = 1.0 / (eps + rx); wY[p] = 1.0 / (eps + ry);
    }
}

// r = b - A x, z = r / diag, p = z; partial sums of r.z and b.b (the latter for the relative tolerance) and, in a second
// pair, r.r
__global__ void __launch_bounds__(L1_BLOCK) k_l1_init(L1Geom g, const double *wU, const double *wX, const double *wY, const double *u, const double *gx, const double *gy,
                                                       const double *x_, double *r, double *z, double *p_, double *partial, double *partial2) {
    double s_rz = 0, s_bb = 0, s_rr = 0;
    for (long long e = blockIdx.x * (long long)L1_BLOCK + threadIdx.x; e < g.n; e += (long long)gridDim.x * L1_BLOCK) {
        const long long p = e / 3;
        const int x = (int)(p % g.w), y = (int)(p / g.w);
        const Arms a = arms_of(g, x, y, p, wU, wX, wY);
        double b = g.alpha * wU[p] * u[e];
        if (x >= 1) b += a.l * gx[e];
        if (x < g.w - 1) b -= a.r * gx[e + 3];
        if (y >= 1) b += a.u * gy[e];
        if (y < g.h - 1) b -= a.d * gy[e + 3LL * g.w];
        const double res = b - apply_at(g, a, e, x, y, x_);
        const double zz = res / a.diag;
        r[e] = res; z[e] = zz; p_[e] = zz;
        s_rz += res * zz; s_bb += b * b; s_rr += res * res;
    }
    block_reduce2(s_rz, s_bb, partial);
    block_reduce2(s_rr, 0.0, partial2);
}

__global__ void __launch_bounds__(L1_BLOCK) k_l1_apply(L1Geom g, const L1Scalars *sc, const double *wU, const double *wX, const double *wY, const double *p_, double *q, double *partial) {
    if (sc->done) return;
    double s = 0;
    for (long long e = blockIdx.x * (long long)L1_BLOCK + threadIdx.x; e < g.n; e += (long long)gridDim.x * L1_BLOCK) {
        const long long p = e / 3;
        const int x = (int)(p % g.w), y = (int)(p / g.w);
        const Arms a = arms_of(g, x, y, p, wU, wX, wY);
        const double v = apply_at(g, a, e, x, y, p_);
        q[e] = v; s += p_[e] * v;
    }
    block_reduce2(s, 0.0, partial);
}

__global__ void __launch_bounds__(L1_BLOCK) k_l1_update(L1Geom g, const L1Scalars *sc, const double *wU, const double *wX, const double *wY, const double *p_, const double *q,
                                                         double *x_, double *r, double *z, double *partial) {
    if (sc->done) return;
    const double step = sc->step;
    double s_rz = 0, s_rr = 0;
    for (long long e = blockIdx.x * (long long)L1_BLOCK + threadIdx.x; e < g.n; e += (long long)gridDim.x * L1_BLOCK) {
        const long long p = e / 3;
        const int x = (int)(p % g.w), y = (int)(p / g.w);
        const Arms a = arms_of(g, x, y, p, wU, wX, wY);
        x_[e] += step * p_[e];
        const double res = r[e] - step * q[e];
        const double zz = res / a.diag;
        r[e] = res; z[e] = zz;
        s_rz += res * zz; s_rr += res * res;
    }
    block_reduce2(s_rz, s_rr, partial);
}

__global__ void __launch_bounds__(L1_BLOCK) k_l1_direction(L1Geom g, const L1Scalars *sc, const double *z, double *p_) {
    if (sc->done) return;
    const double beta = sc->beta;
    for (long long e = blockIdx.x * (long long)L1_BLOCK + threadIdx.x; e < g.n; e += (long long)gridDim.x * L1_BLOCK) p_[e] = z[e] + beta * p_[e];
}

// the scalar step of the recurrences, one thread; `what`: 0 mean|u| -> eps, 1 after init, 2 after apply, 3 after update
__global__ void k_l1_scalars(int what, int nblocks, const double *partial, const double *partial2, L1Scalars *sc, double n, double epsilon, double tol) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double a = 0, b = 0;
    for (int k = 0; k < nblocks; k++) { a += partial[2 * k]; b += partial[2 * k + 1]; }
    if (what == 0) { sc->mean_abs_u = a / n; sc->eps = epsilon * fmax(sc->mean_abs_u, 1e-300); sc->iterations = 0; return; }
    if (what == 1) {
        double rr = 0; for (int k = 0; k < nblocks; k++) rr += partial2[2 * k];
        sc->rz = a; sc->bb = b; sc->rr = rr;
        sc->done = !(rr > tol * tol * b) || !(a > 0.0);
        return;
    }
    if (sc->done) return;
    if (what == 2) { sc->pq = a; if (a > 0.0) sc->step = sc->rz / a; else { sc->step = 0.0; sc->done = 1; } return; }
    // what == 3
    sc->beta = a / sc->rz; sc->rz = a; sc->rr = b; sc->iterations++;
    if (!(b > tol * tol * sc->bb) || !(a > 0.0)) sc->done = 1;
}

struct L1Buffers { DevBuf wU, wX, wY, r, z, p, q, partial, partial2, sc; };

int run_l1(gdpt_ctx *c, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, const gdpt_l1_params *prm, double *d_out, double *info) {
    if (!c || !d_u || !d_gx || !d_gy || !d_out || !prm) return fail(GDPT_ERR_ARG, "null argument");
    if (w < 2 || h < 2) return fail(GDPT_ERR_ARG, "reconstruction needs at least a 2x2 image");
    if (prm->irls_iterations < 0 || prm->irls_iterations > 1000 || prm->cg_iterations < 1 || prm->cg_iterations > 100000)
        return fail(GDPT_ERR_ARG, "gdpt_poisson_l1: irls_iterations outside 0..1000 or cg_iterations outside 1..100000");
    if (prm->start < 0 || prm->start > 2) return fail(GDPT_ERR_ARG, "gdpt_poisson_l1: unknown start %d", prm->start);
    const double alpha = prm->alpha > 0 ? prm->alpha : 0.04, epsilon = prm->epsilon > 0 ? prm->epsilon : 0.05, tol = prm->cg_tolerance > 0 ? prm->cg_tolerance : 1e-6;
    GD_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    L1Geom g; g.w = w; g.h = h; g.n = (long long)w * h * 3; g.alpha = alpha;
    const size_t img = (size_t)g.n * sizeof(double), plane = (size_t)w * h * sizeof(double);
    L1Buffers B;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    auto release = [&]() { DevBuf *all[] = {&B.wU, &B.wX, &B.wY, &B.r, &B.z, &B.p, &B.q, &B.partial, &B.partial2, &B.sc}; for (auto *b : all) b->release(); };
    auto body = [&]() -> int {
        { DevBuf *planes[] = {&B.wU, &B.wX, &B.wY}; for (auto *b : planes) if (int r = b->ensure(plane)) return r; }
        { DevBuf *imgs[] = {&B.r, &B.z, &B.p, &B.q}; for (auto *b : imgs) if (int r = b->ensure(img)) return r; }
        if (int r = B.partial.ensure(L1_GRID * 2 * sizeof(double))) return r;
        if (int r = B.partial2.ensure(L1_GRID * 2 * sizeof(double))) return r;
        if (int r = B.sc.ensure(sizeof(L1Scalars))) return r;
        L1Scalars *sc = B.sc.as<L1Scalars>();
        double *part = B.partial.as<double>(), *part2 = B.partial2.as<double>();
        GD_CUDA(cudaMemsetAsync(sc, 0, sizeof(L1Scalars), st));
        GD_CUDA(cudaEventCreate(&e0)); GD_CUDA(cudaEventCreate(&e1));
        GD_CUDA(cudaEventRecord(e0, st));
        // the starting image
        if (prm->start == GDPT_L1_START_FOURIER) { if (int r = poisson_device(c, w, h, d_u, d_gx, d_gy, alpha, d_out)) return r; }
        else if (prm->start == GDPT_L1_START_PRIMAL) GD_CUDA(cudaMemcpyAsync(d_out, d_u, img, cudaMemcpyDeviceToDevice, st));
        k_l1_mean_abs<<<L1_GRID, L1_BLOCK, 0, st>>>(g, d_u, part);
        k_l1_scalars<<<1, 32, 0, st>>>(0, L1_GRID, part, part2, sc, (double)g.n, epsilon, tol);
        long long launches = 2;
        const int rounds = prm->irls_iterations == 0 ? 1 : prm->irls_iterations;
        for (int it = 0; it < rounds; it++) {
            k_l1_weights<<<L1_GRID, L1_BLOCK, 0, st>>>(g, prm->irls_iterations == 0 ? 1 : 0, prm->reweight_primal != 0, sc, d_out, d_u, d_gx, d_gy, B.wU.as<double>(), B.wX.as<double>(), B.wY.as<double>());
            k_l1_init<<<L1_GRID, L1_BLOCK, 0, st>>>(g, B.wU.as<double>(), B.wX.as<double>(), B.wY.as<double>(), d_u, d_gx, d_gy, d_out,
                                                    B.r.as<double>(), B.z.as<double>(), B.p.as<double>(), part, part2);
            k_l1_scalars<<<1, 32, 0, st>>>(1, L1_GRID, part, part2, sc, (double)g.n, epsilon, tol);
            for (int k = 0; k < prm->cg_iterations; k++) {
                k_l1_apply<<<L1_GRID, L1_BLOCK, 0, st>>>(g, sc, B.wU.as<double>(), B.wX.as<double>(), B.wY.as<double>(), B.p.as<double>(), B.q.as<double>(), part);
                k_l1_scalars<<<1, 32, 0, st>>>(2, L1_GRID, part, part2, sc, (double)g.n, epsilon, tol);
                k_l1_update<<<L1_GRID, L1_BLOCK, 0, st>>>(g, sc, B.wU.as<double>(), B.wX.as<double>(), B.wY.as<double>(), B.p.as<double>(), B.q.as<double>(), d_out,
                                                          B.r.as<double>(), B.z.as<double>(), part);
                k_l1_scalars<<<1, 32, 0, st>>>(3, L1_GRID, part, part2, sc, (double)g.n, epsilon, tol);
                k_l1_direction<<<L1_GRID, L1_BLOCK, 0, st>>>(g, sc, B.z.as<double>(), B.p.as<double>());
            }
            launches += 3 + 5LL * prm->cg_iterations;
        }
        GD_CUDA(cudaGetLastError());
        GD_CUDA(cudaEventRecord(e1, st));
        L1Scalars hs;
        GD_CUDA(cudaMemcpyAsync(&hs, sc, sizeof(hs), cudaMemcpyDeviceToHost, st));
        GD_CUDA(cudaStreamSynchronize(st));
        float ms = 0.f; cudaEventElapsedTime(&ms, e0, e1);
        c->stats.kernel_launches += launches;
        if (info) { info[0] = hs.bb > 0 ? sqrt(hs.rr / hs.bb) : 0.0; info[1] = (double)hs.iterations; info[2] = ms; info[3] = hs.eps; }
        return 0;
    };
    const int rc = body();
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    release();
    return rc;
}

}  // namespace

extern "C" int gdpt_poisson_l1_device(gdpt_ctx *ctx, int w, int h, const double *d_u, const double *d_gx, const double *d_gy, const gdpt_l1_params *params,
                                      double *d_out, double *info) {
    return run_l1(ctx, w, h, d_u, d_gx, d_gy, params, d_out, info);
}

extern "C" int gdpt_poisson_l1(int device, int w, int h, const double *u, const double *gx, const double *gy, const gdpt_l1_params *params, double *out, double *info) {
    if (!u || !gx || !gy || !out || !params) return fail(GDPT_ERR_ARG, "null argument");
    if (w < 2 || h < 2) return fail(GDPT_ERR_ARG, "reconstruction needs at least a 2x2 image");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return fail(GDPT_ERR_CUDA, "no CUDA device: libgdpt has no CPU fallback");
    if (device < 0 || device >= ndev) return fail(GDPT_ERR_ARG, "device out of range");
    GD_CUDA(cudaSetDevice(device));
    gdpt_ctx *c = new gdpt_ctx();                        // a bare context: stream + Poisson plans
    c->device = device; memset(&c->stats, 0, sizeof(c->stats));
    for (auto &e : c->ev) e = nullptr;
    const size_t bytes = (size_t)w * h * 3 * sizeof(double);
    auto body = [&]() -> int {
        GD_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        c->own_stream = true;
        DevBuf *bufs[] = {&c->img, &c->cx, &c->cy, &c->out};
        for (auto *b : bufs) if (int r = b->ensure(bytes)) return r;
        GD_CUDA(cudaMemcpyAsync(c->img.p, u, bytes, cudaMemcpyHostToDevice, c->stream));
        GD_CUDA(cudaMemcpyAsync(c->cx.p, gx, bytes, cudaMemcpyHostToDevice, c->stream));
        GD_CUDA(cudaMemcpyAsync(c->cy.p, gy, bytes, cudaMemcpyHostToDevice, c->stream));
        if (params->start == GDPT_L1_START_OUT) GD_CUDA(cudaMemcpyAsync(c->out.p, out, bytes, cudaMemcpyHostToDevice, c->stream));
        if (int r = run_l1(c, w, h, c->img.as<double>(), c->cx.as<double>(), c->cy.as<double>(), params, c->out.as<double>(), info)) return r;
        GD_CUDA(cudaMemcpyAsync(out, c->out.p, bytes, cudaMemcpyDeviceToHost, c->stream));
        GD_CUDA(cudaStreamSynchronize(c->stream));
        return 0;
    };
    const int rc = body();
    destroy_context(c);
    return rc;
}
