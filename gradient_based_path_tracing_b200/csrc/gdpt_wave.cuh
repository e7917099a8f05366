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

constexpr int TRACE_DEFAULT_VARIANT = 360200;   // one pop site per trip (3.5 %) + 256-bit node loads (2 %) + 12 instead of 20 stack entries in shared memory, i.e. more L1 (1.6 %): profiles/r2_wave_scheduling.txt
constexpr int SORT_MIN_RAYS = 1 << 17;   // below this a launch is too small for the sort to pay

// count_ptr != nullptr: n is only an upper bound, the live count is read on the device (and added to the device-side ray total)
template <bool EXACT, bool COUNT, int V>
inline void launch_trace_v(gdpt_ctx *c, cudaStream_t st, const RayF *rays, HitRec *hits, int n, const int *order, const int *count_ptr) {
    int grid = std::min(blocks_for(n, TRACE_BLOCK), SM_COUNT * 16);
    k_trace<EXACT, COUNT, V><<<grid, TRACE_BLOCK, 0, st>>>(c->bvh, rays, hits, order, n, COUNT ? c->trav_counters.as<unsigned long long>() : nullptr,
                                                            count_ptr, count_ptr ? c->dev_stats.as<unsigned long long>() : nullptr);
}

template <bool EXACT>
inline void launch_trace(gdpt_ctx *c, const RayF *rays, HitRec *hits, int n, const int *order = nullptr, cudaStream_t st = nullptr, const int *count_ptr = nullptr) {
    if (n <= 0) return;
    if (!st) st = c->stream;
    if (c->count_traversal) launch_trace_v<EXACT, true, TRACE_DEFAULT_VARIANT>(c, st, rays, hits, n, order, count_ptr);
    else if constexpr (EXACT) launch_trace_v<EXACT, false, TRACE_DEFAULT_VARIANT>(c, st, rays, hits, n, order, count_ptr);
    else switch (c->trace_variant) {
#define GD_V(v) case v: launch_trace_v<EXACT, false, v>(c, st, rays, hits, n, order, count_ptr); break;
        GD_V(0) GD_V(40000) GD_V(120000) GD_V(200000) GD_V(360000) GD_V(360200) GD_V(360300) GD_V(1000200)
#undef GD_V
        default: launch_trace_v<EXACT, false, TRACE_DEFAULT_VARIANT>(c, st, rays, hits, n, order, count_ptr);
    }
    c->stats.kernel_launches++;
    if (!count_ptr) c->stats.rays_traced += n;
}

template <typename R, int MS, typename... A>
inline void launch_primary(gdpt_ctx *c, cudaStream_t st, int n, A... a) {
    auto go = [&](auto kern) { kern<<<blocks_for(n, 128), 128, 0, st>>>(a...); };
    if constexpr (MS == MS_LAMBERTIAN && sizeof(R) == 4) {   // occupancy variants of the diffuse-only float kernels (GDPT_SHADE_VARIANT)
        switch (c->shade_variant / 10) {
            case 1: go(k_primary<R, MS, 5>); break;
            case 2: go(k_primary<R, MS, 6>); break;
            case 3: go(k_primary<R, MS, 8>); break;
            default: go(k_primary<R, MS, 1>);
        }
    } else go(k_primary<R, MS, 1>);
    c->stats.kernel_launches++;
}

// overflow: the launch was sized by an estimate of the live count; a small grid-stride launch takes the paths beyond it
template <typename R, int MS, typename... A>
inline void launch_bounce(gdpt_ctx *c, cudaStream_t st, int n, bool overflow, const int *order, A... a) {
    if (!c->shade_sorted) order = nullptr;
    auto go = [&](auto kern) { kern<<<blocks_for(n, 128), 128, 0, st>>>(a..., 0, order); };
    if constexpr (MS == MS_LAMBERTIAN && sizeof(R) == 4) {
        switch (c->shade_variant % 10) {
            case 1: go(k_bounce<R, MS, 5>); break;
            case 2: go(k_bounce<R, MS, 6>); break;
            case 3: go(k_bounce<R, MS, 8>); break;
            default: go(k_bounce<R, MS, 1>);
        }
    } else go(k_bounce<R, MS, 1>);
    c->stats.kernel_launches++; c->stats.bounce_iterations++;
    if (overflow) {
        k_bounce<R, MS, 1, true><<<SM_COUNT, 128, 0, st>>>(a..., blocks_for(n, 128) * 128, order);
        c->stats.kernel_launches++;
    }
}

// Buffers and sort parameters of one wave slot.
template <typename R>
int prepare_slot(gdpt_ctx *c, WaveSlot &sl, WaveParams &wp, bool &sorting) {
    const int N = wp.n_paths;
    if (int r = sl.gen->ensure((size_t)N * sizeof(GenRec<R>))) return r;
    if (int r = sl.rays->ensure((size_t)N * 5 * sizeof(RayF))) return r;
    if (int r = sl.hits->ensure((size_t)N * 5 * sizeof(HitRec))) return r;
    if (int r = sl.rec_a->ensure((size_t)N * sizeof(PathRec<R>))) return r;
    if (int r = sl.rec_b->ensure((size_t)N * sizeof(PathRec<R>))) return r;
    if (int r = sl.off->ensure((size_t)N * sizeof(OffRec<R>))) return r;
    if (int r = sl.counters->ensure(64)) return r;
    wp.res = nullptr;
    if (c->deterministic && !wp.rec_out && wp.img) {
        if (int r = sl.res->ensure((size_t)N * 15 * sizeof(double))) return r;
        wp.res = sl.res->as<double>();
    }
    sorting = c->sort_rays && N >= SORT_MIN_RAYS;
    if (sorting) {
        if (int r = sl.sort_keys->ensure((size_t)N * sizeof(int))) return r;
        if (int r = sl.sort_order->ensure((size_t)N * sizeof(int))) return r;
        if (int r = sl.sort_hist->ensure((size_t)(sort_bins(c->sort_cell_bits, c->sort_dir_bits) + 1024) * sizeof(int))) return r;   // histogram + chunk totals
        wp.sort_keys = sl.sort_keys->as<int>(); wp.sort_hist = sl.sort_hist->as<int>();
        wp.sort_cell_bits = c->sort_cell_bits; wp.sort_dir_bits = c->sort_dir_bits;
        const int cells = 1 << c->sort_cell_bits;
        for (int k = 0; k < 3; k++) {
            float ext = c->box_hi[k] - c->box_lo[k];
            wp.sort_lo[k] = c->box_lo[k]; wp.sort_scale[k] = ext > 0.f ? cells / ext : 0.f;
        }
    } else { wp.sort_keys = nullptr; wp.sort_hist = nullptr; }
    return 0;
}

// generate -> trace -> primary -> (trace -> bounce)*, the host reading the live count after every stage.  This is the
// path of record mode, of the first wave of a configuration (it fills the calibration) and of gdpt_set_option("async_waves", 0);
// it also carries the per-kernel event timing (stats.ms_trace / ms_shade).
template <typename R, int MS>
int run_wave_sync(gdpt_ctx *c, DScene<R> sc, WaveParams wp, WaveCalibration *cal) {
    const int N = wp.n_paths;
    if (N <= 0) return 0;
    WaveSlot &sl = c->slot[0];
    bool sorting;
    if (int r = prepare_slot<R>(c, sl, wp, sorting)) return r;
    int *cnt = sl.counters->as<int>();
    const int n_bins = sorting ? sort_bins(c->sort_cell_bits, c->sort_dir_bits) : 0, n_chunks = std::max(1, n_bins / SORT_CHUNK);
    GenRec<R> *gen = sl.gen->as<GenRec<R>>();
    RayF *rays = sl.rays->as<RayF>(); HitRec *hits = sl.hits->as<HitRec>();
    PathRec<R> *cur = sl.rec_a->as<PathRec<R>>(), *nxt = sl.rec_b->as<PathRec<R>>();
    OffRec<R> *off = sl.off->as<OffRec<R>>();
    cudaStream_t st = c->stream;
    cudaEvent_t e0 = c->ev[4], e1 = c->ev[5], e2 = c->ev[6];
    float ms;

    GD_CUDA(cudaMemsetAsync(cnt, 0, 64, st));
    if (wp.res) GD_CUDA(cudaMemsetAsync(wp.res, 0, (size_t)N * 15 * sizeof(double), st));
    k_generate<R><<<blocks_for(N, 128), 128, 0, st>>>(sc, wp, gen, rays);
    c->stats.kernel_launches++;
    GD_CUDA(cudaEventRecord(e0, st));
    launch_trace<ModeSel<R>::exact>(c, rays, hits, 5 * N);
    GD_CUDA(cudaEventRecord(e1, st));
    if (sorting) GD_CUDA(cudaMemsetAsync(wp.sort_hist, 0, (size_t)n_bins * sizeof(int), st));
    launch_primary<R, MS>(c, st, N, sc, wp, gen, hits, cur, off, rays, cnt);
    GD_CUDA(cudaEventRecord(e2, st));
    GD_CUDA(cudaMemcpyAsync(c->h_counts, cnt, sizeof(int), cudaMemcpyDeviceToHost, st));
    GD_CUDA(cudaStreamSynchronize(st));
    cudaEventElapsedTime(&ms, e0, e1); c->stats.ms_trace += ms;
    cudaEventElapsedTime(&ms, e1, e2); c->stats.ms_shade += ms;
    int count = c->h_counts[0];
    if (cal) cal->counts.assign(1, count);
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
            k_sort_scatter<<<blocks_for(count, 256), 256, 0, st>>>(wp.sort_keys, wp.sort_hist, wp.sort_hist + n_bins, count, sl.sort_order->as<int>());
            c->stats.kernel_launches += 3;
            order = sl.sort_order->as<int>();
        }
        launch_trace<ModeSel<R>::exact>(c, rays, hits, count, order);
        GD_CUDA(cudaEventRecord(e1, st));
        if (sorting) GD_CUDA(cudaMemsetAsync(wp.sort_hist, 0, (size_t)n_bins * sizeof(int), st));
        int *cin = cnt + which, *cout = cnt + (which ^ 1);
        GD_CUDA(cudaMemsetAsync(cout, 0, sizeof(int), st));
        launch_bounce<R, MS>(c, st, count, false, order, sc, wp, cur, first ? off : nullptr, hits, cin, count, nxt, rays, cout);
        GD_CUDA(cudaEventRecord(e2, st));
        GD_CUDA(cudaMemcpyAsync(c->h_counts, cout, sizeof(int), cudaMemcpyDeviceToHost, st));
        GD_CUDA(cudaStreamSynchronize(st));
        cudaEventElapsedTime(&ms, e0, e1); c->stats.ms_trace += ms;
        cudaEventElapsedTime(&ms, e1, e2); c->stats.ms_shade += ms;
        count = c->h_counts[0];
        if (cal) cal->counts.push_back(count);
        std::swap(cur, nxt); which ^= 1; first = false;
    }
    if (wp.res) {
        k_wave_reduce<<<blocks_for(wp.npix_pad, 256), 256, 0, st>>>(wp, wp.res, wp.img, wp.cx0, wp.cy0, wp.cx1, wp.cy1);
        c->stats.kernel_launches++;
    }
    GD_CUDA(cudaGetLastError());
    return 0;
}

// The same wave launched from the calibration: every launch is sized by an upper bound of the live count (the calibrated
// count of that stage plus a margin; the kernels are grid-stride and read the true count on the device), the switch to the
// tail kernel happens at the stage where the calibration wave switched, and nothing is read back.  The tail kernel walks
// whatever is still alive to its end, so the schedule never has to be exact.
template <typename R, int MS>
int run_wave_async(gdpt_ctx *c, DScene<R> sc, WaveParams wp, const WaveCalibration &cal, WaveSlot &sl, cudaStream_t st) {
    const int N = wp.n_paths;
    if (N <= 0) return 0;
    bool sorting;
    if (int r = prepare_slot<R>(c, sl, wp, sorting)) return r;
    int *cnt = sl.counters->as<int>();
    const int n_bins = sorting ? sort_bins(c->sort_cell_bits, c->sort_dir_bits) : 0, n_chunks = std::max(1, n_bins / SORT_CHUNK);
    GenRec<R> *gen = sl.gen->as<GenRec<R>>();
    RayF *rays = sl.rays->as<RayF>(); HitRec *hits = sl.hits->as<HitRec>();
    PathRec<R> *cur = sl.rec_a->as<PathRec<R>>(), *nxt = sl.rec_b->as<PathRec<R>>();
    OffRec<R> *off = sl.off->as<OffRec<R>>();
    const double scale = (double)N / (double)cal.n_paths;                 // a shorter last wave
    auto bound = [&](int stage) {
        double b = (stage < (int)cal.counts.size() ? cal.counts[stage] : cal.counts.back()) * scale;
        return (int)std::min<double>(N, b * 1.06 + 16384.0);
    };
    if (wp.res) {
        if (sl.busy) GD_CUDA(cudaStreamWaitEvent(st, sl.reduced, 0));      // the slot's previous wave must have been added to the images
        GD_CUDA(cudaMemsetAsync(wp.res, 0, (size_t)N * 15 * sizeof(double), st));
    }
    GD_CUDA(cudaMemsetAsync(cnt, 0, 64, st));
    k_generate<R><<<blocks_for(N, 128), 128, 0, st>>>(sc, wp, gen, rays);
    c->stats.kernel_launches++;
    launch_trace<ModeSel<R>::exact>(c, rays, hits, 5 * N, nullptr, st);
    if (sorting) GD_CUDA(cudaMemsetAsync(wp.sort_hist, 0, (size_t)n_bins * sizeof(int), st));
    launch_primary<R, MS>(c, st, N, sc, wp, gen, hits, cur, off, rays, cnt);
    c->stats.paths += N;
    int which = 0; bool first = true;
    for (int stage = 0;; stage++) {
        const int est = stage < (int)cal.counts.size() ? (int)(cal.counts[stage] * scale) : 0;     // expected live paths entering this bounce
        const int ub = bound(stage);
        int *cin = cnt + which, *cout = cnt + (which ^ 1);
        if (est <= c->tail_paths) {
            k_tail<R, MS, ModeSel<R>::exact><<<std::max(1, blocks_for(ub, TRACE_BLOCK)), TRACE_BLOCK, 0, st>>>(sc, wp, c->bvh, cur, first ? off : nullptr, rays, N, cin);
            c->stats.kernel_launches++; c->stats.tail_paths += est;
            break;
        }
        const int *order = nullptr;
        if (sorting && est >= SORT_MIN_RAYS) {
            k_sort_scan<<<n_chunks, 1024, 0, st>>>(wp.sort_hist, wp.sort_hist + n_bins);
            k_sort_chunks<<<1, 1024, 0, st>>>(wp.sort_hist + n_bins, n_chunks);
            k_sort_scatter<<<blocks_for(ub, 256), 256, 0, st>>>(wp.sort_keys, wp.sort_hist, wp.sort_hist + n_bins, N, sl.sort_order->as<int>(), cin);
            c->stats.kernel_launches += 3;
            order = sl.sort_order->as<int>();
        }
        launch_trace<ModeSel<R>::exact>(c, rays, hits, N, order, st, cin);
        if (sorting) GD_CUDA(cudaMemsetAsync(wp.sort_hist, 0, (size_t)n_bins * sizeof(int), st));
        GD_CUDA(cudaMemsetAsync(cout, 0, sizeof(int), st));
        launch_bounce<R, MS>(c, st, ub, true, order, sc, wp, cur, first ? off : nullptr, hits, cin, N, nxt, rays, cout);
        std::swap(cur, nxt); which ^= 1; first = false;
    }
    if (wp.res) {     // the reduction of this wave: on the reduce stream, i.e. after the reductions of all earlier waves
        if (!c->reduce_forked) {      // (joins what is already queued on the context's stream: image clears, the calibration wave)
            GD_CUDA(cudaEventRecord(c->ev_fork, c->stream));
            GD_CUDA(cudaStreamWaitEvent(c->reduce_stream, c->ev_fork, 0));
            c->reduce_forked = true;
        }
        GD_CUDA(cudaEventRecord(sl.done, st));
        GD_CUDA(cudaStreamWaitEvent(c->reduce_stream, sl.done, 0));
        k_wave_reduce<<<blocks_for(wp.npix_pad, 256), 256, 0, c->reduce_stream>>>(wp, wp.res, wp.img, wp.cx0, wp.cy0, wp.cx1, wp.cy1);
        GD_CUDA(cudaEventRecord(sl.reduced, c->reduce_stream));
        c->stats.kernel_launches++;
    }
    sl.busy = true;
    GD_CUDA(cudaGetLastError());
    return 0;
}

template <typename R, int MS>
int run_wave_ms(gdpt_ctx *c, DScene<R> sc, WaveParams wp) {
    // record mode, a disabled tail kernel (the loop must then run until the host sees zero) and the counting pass stay synchronous
    const bool can_async = c->async_waves && !wp.xs && c->tail_paths > 0 && !c->count_traversal && !c->time_kernels;
    if (!can_async) return run_wave_sync<R, MS>(c, sc, wp, nullptr);
    WaveCalibration &cal = c->cal;
    const int prec = (int)sizeof(R);
    const bool same = cal.valid && cal.max_depth == sc.max_depth && cal.rr_depth == sc.rr_depth && cal.precision == prec && cal.width == wp.width &&
                      cal.y0 == wp.y0 && cal.y1 == wp.y1 && cal.tail_paths == c->tail_paths && wp.n_paths <= cal.n_paths &&
                      cal.strip_count == wp.strip_count && cal.strip_phase == wp.strip_phase;
    if (!same) {
        cal.valid = false;
        if (int r = run_wave_sync<R, MS>(c, sc, wp, &cal)) return r;
        cal.n_paths = wp.n_paths; cal.max_depth = sc.max_depth; cal.rr_depth = sc.rr_depth; cal.precision = prec; cal.width = wp.width;
        cal.y0 = wp.y0; cal.y1 = wp.y1; cal.tail_paths = c->tail_paths; cal.strip_count = wp.strip_count; cal.strip_phase = wp.strip_phase;
        cal.valid = !cal.counts.empty();
        return 0;
    }
    const int k = c->wave_index % std::max(1, c->wave_streams);
    c->wave_index++;
    WaveSlot &sl = c->slot[k];
    cudaStream_t st = k == 0 ? c->stream : sl.stream;
    if (k > 0 && !(c->forked >> k & 1)) {   // the slot's stream joins the work already queued on the context's stream (image clears, earlier passes)
        GD_CUDA(cudaEventRecord(c->ev_fork, c->stream));
        GD_CUDA(cudaStreamWaitEvent(sl.stream, c->ev_fork, 0));
        c->forked |= 1u << k;
    }
    return run_wave_async<R, MS>(c, sc, wp, cal, sl, st);
}

template <typename R>
int run_wave(gdpt_ctx *c, DScene<R> sc, WaveParams wp) {
    return c->lambertian_only ? run_wave_ms<R, MS_LAMBERTIAN>(c, sc, wp) : run_wave_ms<R, MS_GENERIC>(c, sc, wp);
}

} // namespace gdpt
