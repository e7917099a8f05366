// The wavefront loop for one wave of pixel-samples (host orchestration of the kernels in
// gdpt_wavefront.cuh / gdpt_shade.cuh), templated on the arithmetic type and explicitly
// instantiated in gdpt_wave_f64.cu / gdpt_wave_f32.cu so the two modes compile in parallel.
#pragma once
#include "gdpt_context.cuh"
#include "gdpt_shade.cuh"

namespace gdpt {

constexpr int SM_COUNT = 148;
inline int blocks_for(long long n, int block) { return (int)((n + block - 1) / block); }

template <typename R> struct ModeSel;
template <> struct ModeSel<double> { static constexpr bool exact = true; };
template <> struct ModeSel<float> { static constexpr bool exact = false; };

template <bool EXACT>
inline void launch_trace(gdpt_ctx *c, const RayF *rays, Hit *hits, int n) {
    if (n <= 0) return;
    int grid = std::min(blocks_for(n, TRACE_BLOCK), SM_COUNT * 8);   // persistent blocks, ~7 resident per SM
    if (c->count_traversal) k_trace<EXACT, true><<<grid, TRACE_BLOCK, 0, c->stream>>>(c->bvh, rays, hits, nullptr, n, c->trav_counters.as<unsigned long long>());
    else k_trace<EXACT, false><<<grid, TRACE_BLOCK, 0, c->stream>>>(c->bvh, rays, hits, nullptr, n, nullptr);
    c->stats.kernel_launches++; c->stats.rays_traced += n;
}

// generate -> trace -> primary -> (trace -> bounce)*
template <typename R, int MS>
int run_wave_ms(gdpt_ctx *c, DScene<R> sc, WaveParams wp) {
    const int N = wp.n_paths;
    if (N <= 0) return 0;
    if (int r = c->gen.ensure((size_t)N * sizeof(GenRec<R>))) return r;
    if (int r = c->rays.ensure((size_t)N * 5 * sizeof(RayF))) return r;
    if (int r = c->hits.ensure((size_t)N * 5 * sizeof(Hit))) return r;
    if (int r = c->rec_a.ensure((size_t)N * sizeof(PathRec<R>))) return r;
    if (int r = c->rec_b.ensure((size_t)N * sizeof(PathRec<R>))) return r;
    if (int r = c->off.ensure((size_t)N * sizeof(OffRec<R>))) return r;
    if (int r = c->counters.ensure(64)) return r;
    int *cnt = c->counters.as<int>();
    GenRec<R> *gen = c->gen.as<GenRec<R>>();
    RayF *rays = c->rays.as<RayF>(); Hit *hits = c->hits.as<Hit>();
    PathRec<R> *cur = c->rec_a.as<PathRec<R>>(), *nxt = c->rec_b.as<PathRec<R>>();
    OffRec<R> *off = c->off.as<OffRec<R>>();
    cudaStream_t st = c->stream;
    cudaEvent_t e0 = c->ev[4], e1 = c->ev[5], e2 = c->ev[6];
    float ms;

    GD_CUDA(cudaMemsetAsync(cnt, 0, 64, st));
    k_generate<R><<<blocks_for(N, 128), 128, 0, st>>>(sc, wp, gen, rays);
    c->stats.kernel_launches++;
    GD_CUDA(cudaEventRecord(e0, st));
    launch_trace<ModeSel<R>::exact>(c, rays, hits, 5 * N);
    GD_CUDA(cudaEventRecord(e1, st));
    k_primary<R, MS><<<blocks_for(N, 128), 128, 0, st>>>(sc, wp, gen, hits, cur, off, rays, cnt);
    c->stats.kernel_launches++;
    GD_CUDA(cudaEventRecord(e2, st));
    GD_CUDA(cudaMemcpyAsync(c->h_counts, cnt, sizeof(int), cudaMemcpyDeviceToHost, st));
    GD_CUDA(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ms, e0, e1); c->stats.ms_trace += ms;
    cudaEventElapsedTime(&ms, e1, e2); c->stats.ms_shade += ms;
    int count = c->h_counts[0];
    c->stats.paths += N;
    int which = 0; bool first = true;
    while (count > 0) {
        GD_CUDA(cudaEventRecord(e0, st));
        launch_trace<ModeSel<R>::exact>(c, rays, hits, count);
        GD_CUDA(cudaEventRecord(e1, st));
        int *cin = cnt + which, *cout = cnt + (which ^ 1);
        GD_CUDA(cudaMemsetAsync(cout, 0, sizeof(int), st));
        k_bounce<R, MS><<<blocks_for(count, 128), 128, 0, st>>>(sc, wp, cur, first ? off : nullptr, hits, cin, count, nxt, rays, cout);
        c->stats.kernel_launches++; c->stats.bounce_iterations++;
        GD_CUDA(cudaEventRecord(e2, st));
        GD_CUDA(cudaMemcpyAsync(c->h_counts, cout, sizeof(int), cudaMemcpyDeviceToHost, st));
        GD_CUDA(cudaStreamSynchronize(st));
        cudaEventElapsedTime(&ms, e0, e1); c->stats.ms_trace += ms;
        cudaEventElapsedTime(&ms, e1, e2); c->stats.ms_shade += ms;
        count = c->h_counts[0];
        std::swap(cur, nxt); which ^= 1; first = false;
    }
    GD_CUDA(cudaGetLastError());
    return 0;
}

template <typename R>
int run_wave(gdpt_ctx *c, DScene<R> sc, WaveParams wp) {
    return c->lambertian_only ? run_wave_ms<R, MS_LAMBERTIAN>(c, sc, wp) : run_wave_ms<R, MS_GENERIC>(c, sc, wp);
}

} // namespace gdpt
