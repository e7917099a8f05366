// gdpt_poisson_device_noalloc — the screened-Poisson solve (fourierSolve, /root/reference/src/render.cpp:172-254) for
// callers that may not allocate device memory: code under CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc.  Same kernels and launch sequence as gdpt_poisson_device (gdpt_poisson.cu), but every work
// buffer lives in ONE file-scope __device__ array of fixed size, so the entry point calls no cudaMalloc / cudaFree
// and no allocating library routine; it has no counterpart in the
// reference.
//
//  * Limit: w <= GDPT_NOALLOC_MAX_DIM and h <= GDPT_NOALLOC_MAX_DIM (256): the array holds two planar [3][h][w]
//    work images, the two DCT matrices, the per-block DC partial sums and the DC sums = 4.03 MiB.  Every GPU that
//    loads libgdpt.so holds this array until the process exits.
//  * The array's name is a host-side stand-in in host code, not a device address: its device address is taken
//    once per device, with that device current, by cudaGetSymbolAddress (checked), and that pointer is what the
//    kernels receive.
//  * All calls share the array and it keeps the last call's contents.  Nothing in it has to start at zero: every
//    call rewrites the DCT matrices for its own size and every element a kernel reads was written earlier in the
//    same call (k_poisson_rhs writes all gridDim partial sums that k_dc_reduce reads; the GEMMs write the whole
//    [3][h][w] planes they read back).  Calls that overlap on the GPU (different streams, or two host threads
//    on the same device) RACE on the array: serialise them.
#include "gdpt_context.cuh"
#include "gdpt_poisson.cuh"

namespace {

constexpr int MAXD = 256;                                            // GDPT_NOALLOC_MAX_DIM in include/gdpt.h
constexpr int NB = 148 * 8;                                          // blocks of k_poisson_rhs (as gdpt_poisson.cu)
constexpr size_t PLANE_BYTES = (size_t)3 * MAXD * MAXD * sizeof(double);
constexpr size_t MAT_BYTES = (size_t)MAXD * MAXD * sizeof(double);
constexpr size_t PARTIAL_BYTES = ((size_t)NB * 3 * sizeof(double) + 255) / 256 * 256;
constexpr size_t OFF_A = 0, OFF_B = OFF_A + PLANE_BYTES, OFF_DW = OFF_B + PLANE_BYTES, OFF_DHT = OFF_DW + MAT_BYTES,
                 OFF_PARTIAL = OFF_DHT + MAT_BYTES, OFF_DC = OFF_PARTIAL + PARTIAL_BYTES, SCRATCH_BYTES = OFF_DC + 256;

static __device__ __align__(256) unsigned char g_noalloc_scratch[SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
unsigned char *g_scratch_ptr[MAX_DEVICES] = {nullptr};               // device address of the array, per device

}  // namespace

extern "C" int gdpt_poisson_device_noalloc(int device, int w, int h, const double *d_u, const double *d_gx, const double *d_gy,
                                           double alpha, double *d_out, void *cuda_stream) {
    using namespace gdpt;
    if (!d_u || !d_gx || !d_gy || !d_out) return fail(GDPT_ERR_ARG, "null device image");
    if (w < 2 || h < 2) return fail(GDPT_ERR_ARG, "Poisson solve needs at least a 2x2 image");
    if (w > MAXD || h > MAXD)
        return fail(GDPT_ERR_ARG, "gdpt_poisson_device_noalloc: image %dx%d exceeds the fixed scratch (at most %d x %d); use gdpt_poisson_device", w, h, MAXD, MAXD);
    if (device < 0 || device >= MAX_DEVICES) return fail(GDPT_ERR_ARG, "device %d out of range", device);
    GD_CUDA(cudaSetDevice(device));
    if (!g_scratch_ptr[device]) {                                     // set-up, once per device, with that device current
        void *p = nullptr;
        GD_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return fail(GDPT_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
        g_scratch_ptr[device] = (unsigned char *)p;
    }
    unsigned char *base = g_scratch_ptr[device];
    double *A = (double *)(base + OFF_A), *B = (double *)(base + OFF_B), *Dw = (double *)(base + OFF_DW), *DhT = (double *)(base + OFF_DHT),
           *partial = (double *)(base + OFF_PARTIAL), *dc = (double *)(base + OFF_DC);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (!(alpha > 0)) alpha = 0.04;                                   // render.cpp:353
    k_dct_matrix<<<148 * 4, 256, 0, st>>>(w, 0, Dw);
    k_dct_matrix<<<148 * 4, 256, 0, st>>>(h, 1, DhT);
    k_poisson_rhs<<<NB, RHS_BLOCK, 0, st>>>(w, h, 0, h, d_u, d_gx, d_gy, alpha, A, partial);
    k_dc_reduce<<<1, 96, 0, st>>>(NB, partial, dc);
    Epilogue ep; memset(&ep, 0, sizeof(ep));
    ep.img_w = w; ep.img_h = h; ep.alpha = alpha; ep.dc = dc; ep.out = d_out;
    ep.scale = 4.0 * (w - 1) * (h - 1);
    const long long plane = (long long)w * h;
    auto blocks = [](int n) { return (n + GT - 1) / GT; };
    dim3 g_rows(blocks(w), blocks(3 * h), 1), g_cols(blocks(w), blocks(h), 3);
    ep.mode = EPI_STORE;                                              // forward: rows, then columns + spectral divide + DC
    k_dgemm<<<g_rows, 256, 0, st>>>(3 * h, w, w, A, w, 0, Dw, w, 0, B, w, 0, ep);
    ep.mode = EPI_DIVIDE;
    k_dgemm<<<g_cols, 256, 0, st>>>(h, w, h, DhT, h, 0, B, w, plane, A, w, plane, ep);
    ep.mode = EPI_STORE;                                              // inverse: columns, then rows + scale + RGB interleave
    k_dgemm<<<g_cols, 256, 0, st>>>(h, w, h, DhT, h, 0, A, w, plane, B, w, plane, ep);
    ep.mode = EPI_SCALE_INTERLEAVE;
    k_dgemm<<<g_rows, 256, 0, st>>>(3 * h, w, w, B, w, 0, Dw, w, 0, A, w, 0, ep);
    GD_CUDA(cudaGetLastError());
    return 0;
}
