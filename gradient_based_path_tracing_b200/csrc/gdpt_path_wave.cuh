// Host orchestration of one wave of the `path` integrator (kernels: gdpt_path.cuh):
//   generate -> trace -> shade<FIRST> -> (trace BSDF rays [sorted] + trace shadow rays [any-hit] -> shade)*
#pragma once
#include "gdpt_wave.cuh"
#include "gdpt_path.cuh"

namespace gdpt {

template <bool EXACT>
inline void launch_trace_any(gdpt_ctx *c, const RayF *rays, HitRec *hits, int n, const int *order) {
    if (n <= 0) return;
    int grid = std::min(blocks_for(n, TRACE_BLOCK), SM_COUNT * 16);
    k_trace<EXACT, false, TRACE_DEFAULT_VARIANT, true><<<grid, TRACE_BLOCK, 0, c->stream>>>(c->bvh, rays, hits, order, n, nullptr);
    c->stats.kernel_launches++; c->stats.rays_traced += n;
}

template <typename R, int MS>
int run_path_wave_ms(gdpt_ctx *c, DScene<R> sc, WaveParams wp) {
    const int N = wp.n_paths;
    if (N <= 0) return 0;
    if (int r = c->rays.ensure((size_t)N * 5 * sizeof(RayF))) return r;          // shared with gradpath (5 rays per path there)
    if (int r = c->hits.ensure((size_t)N * 5 * sizeof(HitRec))) return r;
    if (int r = c->shadow_rays.ensure((size_t)N * sizeof(RayF))) return r;
    if (int r = c->shadow_hits.ensure((size_t)N * sizeof(HitRec))) return r;
    if (int r = c->prec_a.ensure((size_t)N * sizeof(PathState<R>))) return r;
    if (int r = c->prec_b.ensure((size_t)N * sizeof(PathState<R>))) return r;
    if (int r = c->counters.ensure(64)) return r;
    const bool sorting = c->sort_rays && N >= SORT_MIN_RAYS;
    if (sorting) {
        if (int r = c->sort_keys.ensure((size_t)N * sizeof(int))) return r;
        if (int r = c->sort_order.ensure((size_t)N * sizeof(int))) return r;
        if (int r = c->sort_hist.ensure((size_t)(sort_bins(c->sort_cell_bits, c->sort_dir_bits) + 1024) * sizeof(int))) return r;
        wp.sort_keys = c->sort_keys.as<int>(); wp.sort_hist = c->sort_hist.as<int>();
        wp.sort_cell_bits = c->sort_cell_bits; wp.sort_dir_bits = c->sort_dir_bits;
        const int cells = 1 << c->sort_cell_bits;
        for (int k = 0; k < 3; k++) {
            float ext = c->box_hi[k] - c->box_lo[k];
            wp.sort_lo[k] = c->box_lo[k]; wp.sort_scale[k] = ext > 0.f ? cells / ext : 0.f;
        }
    } else { wp.sort_keys = nullptr; wp.sort_hist = nullptr; }
    const int n_bins = sorting ? sort_bins(c->sort_cell_bits, c->sort_dir_bits) : 0, n_chunks = std::max(1, n_bins / SORT_CHUNK);
    int *cnt = c->counters.as<int>();
    RayF *rays = c->rays.as<RayF>(), *shadow = c->shadow_rays.as<RayF>();
    HitRec *hits = c->hits.as<HitRec>(), *shadow_hits = c->shadow_hits.as<HitRec>();
    PathState<R> *cur = c->prec_a.as<PathState<R>>(), *nxt = c->prec_b.as<PathState<R>>();
    cudaStream_t st = c->stream;
    cudaEvent_t e0 = c->ev[4], e1 = c->ev[5], e2 = c->ev[6];
    float ms;

    wp.res = nullptr;
    if (c->deterministic && !wp.rec_out && !wp.xs) {      // image renders: per-path result records + ordered reduction (no atomics)
        if (int r = c->res0.ensure((size_t)N * 3 * sizeof(double))) return r;
        wp.res = c->res0.as<double>();
        GD_CUDA(cudaMemsetAsync(wp.res, 0, (size_t)N * 3 * sizeof(double), st));
    }
    GD_CUDA(cudaMemsetAsync(cnt, 0, 64, st));
    k_path_generate<R><<<blocks_for(N, 128), 128, 0, st>>>(sc, wp, cur, rays);
    c->stats.kernel_launches++;
    GD_CUDA(cudaEventRecord(e0, st));
    launch_trace<ModeSel<R>::exact>(c, rays, hits, N);
    GD_CUDA(cudaEventRecord(e1, st));
    if (sorting) GD_CUDA(cudaMemsetAsync(wp.sort_hist, 0, (size_t)n_bins * sizeof(int), st));
    // a shade launch reads the shadow rays of the iteration that just came back while it writes the next ones into
    // compacted slots, so shadow rays ping-pong between two buffers like the path records do
    if (int r = c->off.ensure((size_t)N * sizeof(RayF))) return r;               // second shadow-ray buffer (gradpath's offset records are not live here)
    RayF *shadow_next = c->off.as<RayF>();
    k_path_shade<R, MS, true><<<blocks_for(N, 128), 128, 0, st>>>(sc, wp, cur, hits, nullptr, nullptr, nullptr, N, nxt, rays, shadow, cnt);
    c->stats.kernel_launches++;
    GD_CUDA(cudaEventRecord(e2, st));
    GD_CUDA(cudaMemcpyAsync(c->h_counts, cnt, sizeof(int), cudaMemcpyDeviceToHost, st));
    GD_CUDA(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ms, e0, e1); c->stats.ms_trace += ms;
    cudaEventElapsedTime(&ms, e1, e2); c->stats.ms_shade += ms;
    int count = c->h_counts[0];
    c->stats.paths += N;
    std::swap(cur, nxt);
    int which = 0;
    while (count > 0) {
        if (count <= c->tail_paths) {   // few paths left: finish them in one launch
            GD_CUDA(cudaEventRecord(e0, st));
            k_path_tail<R, MS, ModeSel<R>::exact><<<blocks_for(count, TRACE_BLOCK), TRACE_BLOCK, 0, st>>>(sc, wp, c->bvh, cur, rays, shadow, count);
            GD_CUDA(cudaEventRecord(e1, st));
            c->stats.kernel_launches++; c->stats.tail_paths += count;
            GD_CUDA(cudaEventSynchronize(e1));
            cudaEventElapsedTime(&ms, e0, e1); c->stats.ms_shade += ms;
            break;
        }
        GD_CUDA(cudaEventRecord(e0, st));
        const int *order = nullptr;
        if (sorting && count >= SORT_MIN_RAYS) {
            k_sort_scan<<<n_chunks, 1024, 0, st>>>(wp.sort_hist, wp.sort_hist + n_bins);
            k_sort_chunks<<<1, 1024, 0, st>>>(wp.sort_hist + n_bins, n_chunks);
            k_sort_scatter<<<blocks_for(count, 256), 256, 0, st>>>(wp.sort_keys, wp.sort_hist, wp.sort_hist + n_bins, count, c->sort_order.as<int>());
            c->stats.kernel_launches += 3;
            order = c->sort_order.as<int>();
        }
        launch_trace<ModeSel<R>::exact>(c, rays, hits, count, order);
        // shadow rays leave from the same vertices as the BSDF rays: the permutation sorted by origin cell serves both
        launch_trace_any<ModeSel<R>::exact>(c, shadow, shadow_hits, count, order);
        GD_CUDA(cudaEventRecord(e1, st));
        if (sorting) GD_CUDA(cudaMemsetAsync(wp.sort_hist, 0, (size_t)n_bins * sizeof(int), st));
        int *cin = cnt + which, *cout = cnt + (which ^ 1);
        GD_CUDA(cudaMemsetAsync(cout, 0, sizeof(int), st));
        k_path_shade<R, MS, false><<<blocks_for(count, 128), 128, 0, st>>>(sc, wp, cur, hits, shadow, shadow_hits, cin, count, nxt, rays, shadow_next, cout,
                                                                           c->shade_sorted ? order : nullptr);
        c->stats.kernel_launches++; c->stats.bounce_iterations++;
        GD_CUDA(cudaEventRecord(e2, st));
        GD_CUDA(cudaMemcpyAsync(c->h_counts, cout, sizeof(int), cudaMemcpyDeviceToHost, st));
        GD_CUDA(cudaStreamSynchronize(st));
        cudaEventElapsedTime(&ms, e0, e1); c->stats.ms_trace += ms;
        cudaEventElapsedTime(&ms, e1, e2); c->stats.ms_shade += ms;
        count = c->h_counts[0];
        std::swap(cur, nxt); std::swap(shadow, shadow_next); which ^= 1;
    }
    if (wp.res) { k_path_reduce<<<blocks_for(wp.npix_pad, 256), 256, 0, st>>>(wp, wp.res, wp.img); c->stats.kernel_launches++; }
    GD_CUDA(cudaGetLastError());
    return 0;
}

template <typename R>
int run_path_wave(gdpt_ctx *c, DScene<R> sc, WaveParams wp) {
    return c->lambertian_only ? run_path_wave_ms<R, MS_LAMBERTIAN>(c, sc, wp) : run_path_wave_ms<R, MS_GENERIC>(c, sc, wp);
}

} // namespace gdpt
