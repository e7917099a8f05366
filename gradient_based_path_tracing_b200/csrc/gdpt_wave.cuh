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

constexpr int TRACE_DEFAULT_VARIANT = 0;
constexpr int SORT_MIN_RAYS = 1 << 17;   // below this a launch is too small for the sort to pay

template <bool EXACT, bool COUNT, int V>
inline void launch_trace_v(gdpt_ctx *c, const RayF *rays, HitRec *hits, int n, const int *order) {
    int grid = std::min(blocks_for(n, TRACE_BLOCK), SM_COUNT * 16);
    k_trace<EXACT, COUNT, V><<<grid, TRACE_BLOCK, 0, c->stream>>>(c->bvh, rays, hits, order, n, COUNT ? c->trav_counters.as<unsigned long long>() : nullptr);
}

template <bool EXACT>
inline void launch_trace(gdpt_ctx *c, const RayF *rays, HitRec *hits, int n, const int *order = nullptr) {
    if (n <= 0) return;
    if (c->count_traversal) launch_trace_v<EXACT, true, TRACE_DEFAULT_VARIANT>(c, rays, hits, n, order);
    else if constexpr (EXACT) launch_trace_v<EXACT, false, TRACE_DEFAULT_VARIANT>(c, rays, hits, n, order);
    else switch (c->trace_variant) {
#define GD_V(v) case v: launch_trace_v<EXACT, false, v>(c, rays, hits, n, order); break;
        GD_V(0) GD_V(10000) GD_V(20000) GD_V(30000)
#undef GD_V
        default: launch_trace_v<EXACT, false, TRACE_DEFAULT_VARIANT>(c, rays, hits, n, order);
    }
    c->stats.kernel_launches++; c->stats.rays_traced += n;
}

// generate -> trace -> primary -> (trace -> bounce)*
template <typename R, int MS>
int run_wave_ms(gdpt_ctx *c, DScene<R> sc, WaveParams wp) {
    const int N = wp.n_paths;
    if (N <= 0) return 0;
    if (int r = c->gen.ensure((size_t)N * sizeof(GenRec<R>))) return r;
    if (int r = c->rays.ensure((size_t)N * 5 * sizeof(RayF))) return r;
    if (int r = c->hits.ensure((size_t)N * 5 * sizeof(HitRec))) return r;
    if (int r = c->rec_a.ensure((size_t)N * sizeof(PathRec<R>))) return r;
    if (int r = c->rec_b.ensure((size_t)N * sizeof(PathRec<R>))) return r;
    if (int r = c->off.ensure((size_t)N * sizeof(OffRec<R>))) return r;
    if (int r = c->counters.ensure(64)) return r;
    const bool sorting = c->sort_rays && N >= SORT_MIN_RAYS;
    if (sorting) {
        if (int r = c->sort_keys.ensure((size_t)N * sizeof(int))) return r;
        if (int r = c->sort_order.ensure((size_t)N * sizeof(int))) return r;
        if (int r = c->sort_hist.ensure((size_t)(sort_bins(c->sort_cell_bits, c->sort_dir_bits) + 1024) * sizeof(int))) return r;   // histogram + chunk totals
        wp.sort_keys = c->sort_keys.as<int>(); wp.sort_hist = c->sort_hist.as<int>();
        wp.sort_cell_bits = c->sort_cell_bits; wp.sort_dir_bits = c->sort_dir_bits;
        const int cells = 1 << c->sort_cell_bits;
        for (int k = 0; k < 3; k++) {
            float ext = c->box_hi[k] - c->box_lo[k];
            wp.sort_lo[k] = c->box_lo[k]; wp.sort_scale[k] = ext > 0.f ? cells / ext : 0.f;
        }
    } else { wp.sort_keys = nullptr; wp.sort_hist = nullptr; }
    int *cnt = c->counters.as<int>();
    const int n_bins = sorting ? sort_bins(c->sort_cell_bits, c->sort_dir_bits) : 0, n_chunks = std::max(1, n_bins / SORT_CHUNK);
    GenRec<R> *gen = c->gen.as<GenRec<R>>();
    RayF *rays = c->rays.as<RayF>(); HitRec *hits = c->hits.as<HitRec>();
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
    if (sorting) GD_CUDA(cudaMemsetAsync(wp.sort_hist, 0, (size_t)n_bins * sizeof(int), st));
    {
        auto args = [&](auto kern) { kern<<<blocks_for(N, 128), 128, 0, st>>>(sc, wp, gen, hits, cur, off, rays, cnt); };
        if constexpr (MS == MS_LAMBERTIAN && sizeof(R) == 4) {   // occupancy variants of the diffuse-only float kernels (GDPT_SHADE_VARIANT)
            switch (c->shade_variant / 10) {
                case 1: args(k_primary<R, MS, 5>); break;
                case 2: args(k_primary<R, MS, 6>); break;
                case 3: args(k_primary<R, MS, 8>); break;
                default: args(k_primary<R, MS, 1>);
            }
        } else args(k_primary<R, MS, 1>);
    }
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
        if (count <= c->tail_paths) {   // few paths left: finish them in one launch
            GD_CUDA(cudaEventRecord(e0, st));
            k_tail<R, MS, ModeSel<R>::exact><<<blocks_for(count, TRACE_BLOCK), TRACE_BLOCK, 0, st>>>(sc, wp, c->bvh, cur, first ? off : nullptr, rays, count);
            GD_CUDA(cudaEventRecord(e1, st));
            c->stats.kernel_launches++; c->stats.tail_paths += count;
            GD_CUDA(cudaEventSynchronize(e1));
            cudaEventElapsedTime(&ms, e0, e1); c->stats.ms_shade += ms;
            break;
        }
        GD_CUDA(cudaEventRecord(e0, st));
        const int *order = nullptr;
        if (sorting && count >= SORT_MIN_RAYS) {   // histogram was filled by the kernel that emitted these rays
            k_sort_scan<<<n_chunks, 1024, 0, st>>>(wp.sort_hist, wp.sort_hist + n_bins);
            k_sort_chunks<<<1, 1024, 0, st>>>(wp.sort_hist + n_bins, n_chunks);
            k_sort_scatter<<<blocks_for(count, 256), 256, 0, st>>>(wp.sort_keys, wp.sort_hist, wp.sort_hist + n_bins, count, c->sort_order.as<int>());
            c->stats.kernel_launches += 3;
            order = c->sort_order.as<int>();
        }
        launch_trace<ModeSel<R>::exact>(c, rays, hits, count, order);
        GD_CUDA(cudaEventRecord(e1, st));
        if (sorting) GD_CUDA(cudaMemsetAsync(wp.sort_hist, 0, (size_t)n_bins * sizeof(int), st));
        int *cin = cnt + which, *cout = cnt + (which ^ 1);
        GD_CUDA(cudaMemsetAsync(cout, 0, sizeof(int), st));
        {
            auto args = [&](auto kern) { kern<<<blocks_for(count, 128), 128, 0, st>>>(sc, wp, cur, first ? off : nullptr, hits, cin, count, nxt, rays, cout); };
            if constexpr (MS == MS_LAMBERTIAN && sizeof(R) == 4) {
                switch (c->shade_variant % 10) {
                    case 1: args(k_bounce<R, MS, 5>); break;
                    case 2: args(k_bounce<R, MS, 6>); break;
                    case 3: args(k_bounce<R, MS, 8>); break;
                    default: args(k_bounce<R, MS, 1>);
                }
            } else args(k_bounce<R, MS, 1>);
        }
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
