// Device-side texture ingestion: gdpt_create uploads only what the description holds — level 0 of
// every image, as doubles or as bytes indexing a palette (8-bit files: 256 decoded values) — and the palette expansion,
// the mip chains (make_mipmap, mipmap.h:28-44) and the float copies of the pools are built by these kernels.  The
// arithmetic is the host builder's, operation for operation ((a + b + c + d) * 0.25 in double, same flat-index clamp),
// so the exact mode's texels are the reference's bit for bit; what changes is where the 70 MB of Sponza texels are
// produced: in HBM instead of on one host core per image followed by a pageable upload.
#include "gdpt_context.cuh"

namespace gdpt {

// level sizes and pool offsets (in texels) of one image; `cursor` = texels already in the pool of its kind
static void mip_layout(int w, int h, int C, DImage &im, long long &cursor) {
    int size = w > h ? w : h;
    int levels = std::min((int)ceil(log2((double)size) + 1), 8);
    im.n_levels = levels; im.channels = C;
    int pw = w, ph = h;
    for (int l = 0; l < levels; l++) {
        if (l > 0) { pw = std::max(pw / 2, 1); ph = std::max(ph / 2, 1); }
        im.w[l] = pw; im.h[l] = ph; im.off[l] = cursor;
        cursor += (long long)pw * ph;
    }
    for (int l = levels; l < 8; l++) { im.w[l] = im.h[l] = 1; im.off[l] = 0; }
}

// level 0 from a palette: dst[i] = palette[index[i]]; an index beyond the palette raises *err
__global__ void __launch_bounds__(256) k_palette_expand(const uint8_t *__restrict__ index, const double *__restrict__ palette, int palette_size,
                                                        long long n, double *__restrict__ dst, int *err) {
    __shared__ double pal[256];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) pal[i] = i < palette_size ? palette[i] : 0.0;
    __syncthreads();
    bool bad = false;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        int b = index[i];
        bad |= b >= palette_size;
        dst[i] = pal[b];
    }
    if (bad) *err = 1;
}

// one mip level: 2x2 box over the previous level.  The reference reads prev(2x+1, 2y+1) through the FLAT index
// y*width + x even when a dimension has collapsed to 1 (it runs past the row, and for the last element past the image);
// the flat index is reproduced and its one out-of-bounds element clamped ("quirks kept").
template <int C>
__global__ void __launch_bounds__(256) k_mip_level(const double *__restrict__ prev, int pw, int ph, double *__restrict__ next, int nw, int nh) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)nw * nh) return;
    const int x = (int)(i % nw), y = (int)(i / nw);
    const long long plast = (long long)pw * ph - 1;
    auto at = [&](int xx, int yy) { long long k = (long long)yy * pw + xx; return k > plast ? plast : k; };
    const long long a = at(2 * x, 2 * y), b = at(2 * x + 1, 2 * y), cc = at(2 * x, 2 * y + 1), dd = at(2 * x + 1, 2 * y + 1);
#pragma unroll
    for (int c = 0; c < C; c++) next[i * C + c] = (prev[a * C + c] + prev[b * C + c] + prev[cc * C + c] + prev[dd * C + c]) * 0.25;
}

__global__ void __launch_bounds__(256) k_texels3_to_float(const double *__restrict__ src, float4 *__restrict__ dst, long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = make_float4((float)src[3 * i], (float)src[3 * i + 1], (float)src[3 * i + 2], 0.f);
}
__global__ void __launch_bounds__(256) k_texels1_to_float(const double *__restrict__ src, float *__restrict__ dst, long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = (float)src[i];
}

static inline int grid_for(long long n, int cap = 148 * 32) { return (int)std::max<long long>(1, std::min<long long>((n + 255) / 256, cap)); }

// Fills im3 / im1 (level tables) and the four texel pools of the context.  Asynchronous on c->stream; the caller
// synchronises once at the end of gdpt_create.
int build_texture_pools(gdpt_ctx *c, const gdpt_scene_desc *d, std::vector<DImage> &im3, std::vector<DImage> &im1) {
    const int n3 = d->n_images3, n1 = d->n_images1;
    im3.assign(n3, DImage()); im1.assign(n1, DImage());
    long long total3 = 0, total1 = 0, stage_bytes = 16;
    for (int i = 0; i < n3 + n1; i++) {
        const gdpt_image &gi = i < n3 ? d->images3[i] : d->images1[i - n3];
        const int C = i < n3 ? 3 : 1;
        if (gi.channels != C || gi.width <= 0 || gi.height <= 0) return fail(GDPT_ERR_ARG, "bad image (%d x %d x %d)", gi.width, gi.height, gi.channels);
        if (!gi.data && !(gi.palette && gi.index)) return fail(GDPT_ERR_ARG, "image has neither data nor palette");
        if (!gi.data && (gi.palette_size < 1 || gi.palette_size > 256)) return fail(GDPT_ERR_ARG, "palette of %d entries (1..256 expected)", gi.palette_size);
        mip_layout(gi.width, gi.height, C, i < n3 ? im3[i] : im1[i - n3], i < n3 ? total3 : total1);
        if (!gi.data) stage_bytes += 256 * (long long)sizeof(double) + (((long long)gi.width * gi.height * C + 255) & ~255LL);
    }
    if (int r = c->texels3_d.ensure(std::max<size_t>((size_t)total3 * 3 * sizeof(double), 16))) return r;
    if (int r = c->texels1_d.ensure(std::max<size_t>((size_t)total1 * sizeof(double), 16))) return r;
    if (int r = c->texels3_f.ensure(std::max<size_t>((size_t)total3 * sizeof(float4), 16))) return r;
    if (int r = c->texels1_f.ensure(std::max<size_t>((size_t)total1 * sizeof(float), 16))) return r;
    DevBuf stage;
    struct Release { DevBuf &b; ~Release() { b.release(); } } release{stage};
    if (int r = stage.ensure((size_t)stage_bytes)) return r;
    cudaStream_t st = c->stream;
    int *err = stage.as<int>();
    GD_CUDA(cudaMemsetAsync(err, 0, 16, st));
    long long cursor = 16;
    for (int i = 0; i < n3 + n1; i++) {
        const gdpt_image &gi = i < n3 ? d->images3[i] : d->images1[i - n3];
        const DImage &im = i < n3 ? im3[i] : im1[i - n3];
        const int C = i < n3 ? 3 : 1;
        double *pool = i < n3 ? c->texels3_d.as<double>() : c->texels1_d.as<double>();
        const long long n = (long long)gi.width * gi.height * C;
        double *lvl0 = pool + im.off[0] * C;
        if (gi.data) GD_CUDA(cudaMemcpyAsync(lvl0, gi.data, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, st));
        else {
            double *pal = (double *)((char *)stage.p + cursor); cursor += 256 * (long long)sizeof(double);
            uint8_t *idx = (uint8_t *)stage.p + cursor; cursor += (n + 255) & ~255LL;
            GD_CUDA(cudaMemcpyAsync(pal, gi.palette, (size_t)gi.palette_size * sizeof(double), cudaMemcpyHostToDevice, st));
            GD_CUDA(cudaMemcpyAsync(idx, gi.index, (size_t)n, cudaMemcpyHostToDevice, st));
            k_palette_expand<<<grid_for(n), 256, 0, st>>>(idx, pal, gi.palette_size, n, lvl0, err);
        }
        for (int l = 1; l < im.n_levels; l++) {
            const double *prev = pool + im.off[l - 1] * C; double *next = pool + im.off[l] * C;
            const long long m = (long long)im.w[l] * im.h[l];
            const int grid = (int)((m + 255) / 256);
            if (C == 3) k_mip_level<3><<<grid, 256, 0, st>>>(prev, im.w[l - 1], im.h[l - 1], next, im.w[l], im.h[l]);
            else k_mip_level<1><<<grid, 256, 0, st>>>(prev, im.w[l - 1], im.h[l - 1], next, im.w[l], im.h[l]);
        }
    }
    if (total3 > 0) k_texels3_to_float<<<(int)((total3 + 255) / 256), 256, 0, st>>>(c->texels3_d.as<double>(), c->texels3_f.as<float4>(), total3);
    if (total1 > 0) k_texels1_to_float<<<(int)((total1 + 255) / 256), 256, 0, st>>>(c->texels1_d.as<double>(), c->texels1_f.as<float>(), total1);
    int h_err = 0;
    GD_CUDA(cudaMemcpyAsync(&h_err, err, sizeof(int), cudaMemcpyDeviceToHost, st));
    GD_CUDA(cudaStreamSynchronize(st));          // the staging buffer is released on return
    GD_CUDA(cudaGetLastError());
    if (h_err) return fail(GDPT_ERR_ARG, "palette index out of range");
    return 0;
}

} // namespace gdpt
