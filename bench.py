#!/usr/bin/env python
"""bench.py — GDPT Mpaths/s (base + 4 shifts) on Sponza 1920x1080, 512 spp (BASELINE.json configs[3]).

A step = one full gradient_path_render of the workload: W*H*spp pixel-samples through the
wavefront (5 primaries + bounce rays each), merge, screened-Poisson solve.
  value  device-resident: scene in HBM, images stay on the device; CUDA events, max over ranks
  e2e    the reference-facing call: gdpt_create (scene upload + BVH) + gdpt_render into HOST
         buffers (4 images copied back) + destroy, every step
  --impl reference: the reference's own CPU gradpath (oracle/_ref = /root/reference/src
         compiled with Embree/FFTW stand-ins) on the box's host cores, bounded sample.
N>1: one process per GPU (torchrun), image split into row bands, one NCCL reduce onto rank 0
(strong scaling: the image is fixed).  Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "GDPT Mpaths/s (base+4 shifts), Sponza 1920x1080"
UNIT = "Mpaths/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--scene", default="sponza_1080p")
    ap.add_argument("--spp", type=int, default=0, help="0 = the config's sample count (512)")
    ap.add_argument("--precision", default="f32", choices=["f32", "f64"])
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="target CPU time of one reference sample")
    ap.add_argument("--ref-total-seconds", type=float, default=150.0, help="--impl reference: target wall time of all warm-up + timed steps together")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--f64-steps", type=int, default=1, help="N=1: full-size steps of the FP64 (reference-precision) mode timed beside the headline (0 = skip)")
    ap.add_argument("--offsets", default="reference", choices=["reference", "shift"],
                    help="gradpath offset semantics: the reference's (the metric's config) or shift-mapped offsets (opt-in mode, DESIGN.md 12)")
    ap.add_argument("--camera", type=int, default=0, help="camera of the sweep: 0 = the scene's own lookat; 1, 2 = synthetic origins inside the Sponza atrium (SURVEY 8d)")
    ap.add_argument("--share-bvh", type=int, default=-1, help="N>1: 1 = rank 0 builds the BVH and broadcasts it, 0 = every rank builds its own, -1 = the library default")
    ap.add_argument("--solver", default="auto", choices=["auto", "gather", "distributed"],
                    help="N>1 Poisson solve: bands summed onto rank 0, or the all-to-all transposed solve (auto: distributed from 4K up)")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                pass
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except (ValueError, IndexError):
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


# --------------------------------------------------------------------------------------- reference arm
def metric_name(w, h):
    return METRIC if (w, h) == (1920, 1080) else METRIC.replace("1920x1080", "%dx%d" % (w, h))


def workload_name(scene, w, h, spp, offsets="reference"):
    """The ONE workload string both arms print (config.workload)."""
    return "%s gradpath %dx%d %d spp, %s" % (scene, w, h, spp, "reference offset semantics" if offsets == "reference" else "shift-mapped offsets (opt-in mode)")


class RefTimer:
    """Times the compiled reference's render() (oracle/_ref) and models its cost as AFFINE in the sample count:
    cost(spp) = fixed + spp * per_spp.  `fixed` is what render.cpp does once per frame on one thread — image
    allocation, the merge (render.cpp:340-350) and fourierSolve (:353) — `per_spp` is the sampling loop (:277-330)
    on all host threads.  Two calibration renders (1 spp and `hi` spp) give the fit; every timed step is one
    render at `spp_step` whose slope is re-derived with the calibrated `fixed`."""

    def __init__(self, scene_name, cores):
        from oracle import ref
        self.ref = ref
        self.cores = cores
        ref.init(cores)
        self.s = ref.RefScene(ref.scene_path(scene_name))
        self.w, self.h, self.full_spp = self.s.width, self.s.height, self.s.spp

    def render(self, spp):
        ref = self.ref
        self.s.set_options(spp=spp)
        ref.rtc_stats(True)
        c0 = ref.tsc(); t0 = time.perf_counter()
        self.s.render()
        dt = time.perf_counter() - t0
        ticks, calls = ref.rtc_stats()
        hz = (ref.tsc() - c0) / dt
        return dt, ticks / hz          # wall seconds, thread-seconds inside rtcIntersect1/rtcOccluded1

    def calibrate(self, budget_s):
        t1, r1 = self.render(1)
        hi = max(2, min(8, int(budget_s / max(t1, 1e-3))))
        th, rh = self.render(hi)
        self.per_spp = max((th - t1) / (hi - 1), 1e-9)
        self.fixed = max(t1 - self.per_spp, 0.0)
        self.rtc_per_spp = max((rh - r1) / (hi - 1), 0.0)      # thread-seconds per spp inside the stand-in ray caster
        return hi

    def solve_alone(self):
        import numpy as np
        u = np.zeros((self.h, self.w, 3)); u[::7, ::5] = 1.0
        t0 = time.perf_counter(); self.ref.fourier_solve(u, u, u); return time.perf_counter() - t0

    def close(self):
        self.s.close()


def reference_line(args, timer, full_spp, step_times, spp_step, hi):
    w, h, cores = timer.w, timer.h, timer.cores
    slopes = [max((t - timer.fixed) / spp_step, 1e-9) for t in step_times] or [timer.per_spp]
    per_spp = sum(slopes) / len(slopes)
    cost = timer.fixed + full_spp * per_spp                     # seconds for the full workload
    mpaths = 5.0 * w * h * full_spp / cost / 1e6
    rtc_frac = min(1.0, timer.rtc_per_spp / (timer.per_spp * cores)) if timer.per_spp > 0 else None
    # the same run with the ray caster free of charge: a bound no Embree build could beat
    cost_free = timer.fixed + full_spp * per_spp * (1.0 - (rtc_frac or 0.0))
    sample = ("%dx%d: render() timed at 1 and %d spp for the affine fit cost(spp) = %.2f s + spp * %.3f s, then %d step(s) at %d spp "
              "(of %d), %d host threads; fourierSolve alone %.2f s; ray casting = oracle stand-in (BVH4, SSE slab tests, double "
              "Moeller-Trumbore; no Embree on this box) = %.0f %% of the sampling loop's thread time" %
              (w, h, hi, timer.fixed, per_spp, len(step_times), spp_step, full_spp, cores, timer.solve_s, 100 * (rtc_frac or 0)))
    cb = {"value": mpaths, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample,
          "fixed_s": timer.fixed, "per_spp_s": per_spp, "solve_alone_s": timer.solve_s, "rtc_fraction": rtc_frac,
          "value_if_ray_casting_were_free": 5.0 * w * h * full_spp / cost_free / 1e6}
    return mpaths, cost, cb


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref
    if not ref.scenes_available():
        emit({"impl": "reference", "unavailable": "oracle/_ref (compiled reference + staged scenes) is missing"})
        return
    cores = os.cpu_count() or 1
    timer = RefTimer(args.scene, cores)
    full_spp = args.spp or timer.full_spp
    hi = timer.calibrate(args.cpu_seconds)
    timer.solve_s = timer.solve_alone()
    # every step is one bounded render; the whole --steps/--warmup run is held to about --ref-total-seconds
    n = max(1, args.warmup + args.steps)
    spp_step = max(1, min(full_spp, int((args.ref_total_seconds / n - timer.fixed) / timer.per_spp)))
    times = []
    for i in range(n):
        dt, _ = timer.render(spp_step)
        if i >= args.warmup:
            times.append(dt)
    mpaths, cost, cb = reference_line(args, timer, full_spp, times, spp_step, hi)
    timer.close()
    line = {"metric": metric_name(timer.w, timer.h), "value": mpaths, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": cost * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "impl": "reference",
            "config": {"workload": workload_name(args.scene, timer.w, timer.h, full_spp)},
            "cpu_baseline": cb,
            "e2e": {"value": mpaths, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def cpu_baseline(args, scene_name, full_spp):
    """The reference on the host cores, bounded sample (rank 0, N=1 only).  Runs in a child process: the reference keeps
    `assert` on in release builds (lajolla.h:3-5) and aborts on some scenes (the Disney config trips
    microfacet.h:36 under gradpath) — that must not take the bench down with it."""
    try:
        cmd = [sys.executable, os.path.abspath(__file__), "--impl", "reference", "--scene", scene_name, "--spp", str(full_spp), "--steps", "1",
               "--warmup", "0", "--cpu-seconds", str(args.cpu_seconds), "--ref-total-seconds", str(max(6.0, args.cpu_seconds))]
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
        if r.returncode != 0:
            tail = [l for l in r.stderr.strip().splitlines() if l.strip()][-1:] or ["no output"]
            return {"value": None, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "reference",
                    "sample": "the reference itself aborted on this scene (exit %d): %s" % (r.returncode, tail[0][-300:])}
        line = json.loads(r.stdout.strip().splitlines()[-1])
        if "cpu_baseline" not in line:
            return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": line.get("unavailable", "no baseline")}
        return line["cpu_baseline"]
    except Exception as e:  # the baseline must never take the bench down
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "failed: %r" % (e,)}


def synthetic_camera(k):
    """cam_to_world of the camera sweep's synthetic views, built like the reference's look_at (src/transform.cpp:46-68).
    Sponza spans x -17.4..17.4, y -0.9..15.7, z -7.8..7.8; the shipped camera sits at (14.4, 6.9, -1.2) looking down -x."""
    import numpy as np
    pos, look = {1: ((0.0, 2.0, 0.0), (14.0, 5.0, 3.0)),          # floor of the atrium, looking up a long side towards the arcades
                 2: ((-12.0, 9.0, 2.5), (10.0, 2.0, -3.0))}[k]     # upper gallery, looking down the length of the hall
    pos, look, up = np.array(pos), np.array(look), np.array((0.0, 1.0, 0.0))
    d = (look - pos) / np.linalg.norm(look - pos)
    left = np.cross(up, d); left /= np.linalg.norm(left)
    new_up = np.cross(d, left)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, pos
    return m


def equal_time_block(g, device, budget_s=1.0, ref_spp=4096):
    """The second half of the metric: relMSE at equal TIME on the shipped Cornell box (BASELINE configs[0], 512x512) — the GPU
    product and the compiled reference each get `budget_s` of wall clock for gradient_path_render; both are judged against the
    same converged image (the GPU's own `ref_spp` render of the same estimator).  relMSE = mean((I - R)^2 / (R^2 + 1e-3))."""
    import numpy as np
    relmse = lambda i, r: float(np.mean((i - r) ** 2 / (r ** 2 + 1e-3)))
    out = {"scene": "cbox_gdpt 512x512", "budget_s": budget_s, "reference_spp": ref_spp}
    try:
        sc = g.Scene.named("cbox_gdpt")
        g.gradient_path_render(sc, device=device, spp=2, precision="f32")
        t0 = time.perf_counter(); R = g.gradient_path_render(sc, device=device, spp=ref_spp, precision="f32", seed=99); tr = time.perf_counter() - t0
        n = max(1, int(ref_spp / tr * budget_s))
        t0 = time.perf_counter(); a = g.gradient_path_render(sc, device=device, spp=n, precision="f32", seed=7); ta = time.perf_counter() - t0
        out["gpu"] = {"spp": n, "seconds": ta, "relmse_primal": relmse(a.primal, R.primal), "relmse_poisson": relmse(a.poisson, R.poisson)}
        sc.close()
        from oracle import ref
        if ref.scenes_available():
            ref.init(os.cpu_count() or 1)
            cs = ref.RefScene(ref.scene_path("cbox_gdpt"))
            cs.set_options(spp=1)
            t0 = time.perf_counter(); cs.render(); t1 = time.perf_counter() - t0
            m = max(1, int(round(budget_s / t1)))
            cs.set_options(spp=m)
            t0 = time.perf_counter(); c = cs.render(); tc = time.perf_counter() - t0
            cs.close()
            out["cpu_reference"] = {"spp": m, "seconds": tc, "cores": os.cpu_count() or 1, "relmse_primal": relmse(c["primal"], R.primal),
                                    "relmse_poisson": relmse(c["poisson"], R.poisson)}
    except Exception as e:
        out["failed"] = repr(e)
    return out


# --------------------------------------------------------------------------------------- native arm
def run_native(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import gradient_based_path_tracing_b200 as g
    from gradient_based_path_tracing_b200 import _capi, multi
    from gradient_based_path_tracing_b200.render import _params

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or _capi.lib().gdpt_device_count() < 1:
        raise RuntimeError("bench.py: no CUDA device; the product has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        class _Solo:
            ReduceOp = type("R", (), {"SUM": 0})
            def get_rank(self): return 0
            def get_world_size(self): return 1
            def reduce(self, *a, **k): pass
            def barrier(self): pass
        dist_ = _Solo()
    dist_ = dist if world > 1 else dist_

    sc = g.Scene.named(args.scene)
    if args.camera:
        sc = g.Scene(dict(sc.a, **{"camera.cam_to_world": synthetic_camera(args.camera)}))
    spp = args.spp or sc.options.samples_per_pixel
    w, h = sc.width, sc.height
    lib, check = _capi.lib(), _capi.check
    share = {} if args.share_bvh < 0 else {"share_bvh": bool(args.share_bvh)}
    R = multi.DistributedRenderer(sc, local, dist_, torch, solver=args.solver, **share)
    check(lib.gdpt_set_stream(R.ctx, C.c_void_p(torch.cuda.current_stream().cuda_stream)))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def stats():
        s = _capi.GdptStats()
        check(lib.gdpt_get_stats(R.ctx, C.byref(s)))
        return {k: getattr(s, k) for k, _ in s._fields_}

    for _ in range(args.warmup):
        R.render(spp, precision=args.precision, offsets=args.offsets)
    barrier()
    s0 = stats()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        R.render(spp, precision=args.precision, offsets=args.offsets)
    e1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    s1 = stats()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    launches = torch.tensor([s1["kernel_launches"] - s0["kernel_launches"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(launches, op=dist.ReduceOp.SUM)
    ms_step = float(ms.item()) / args.steps
    paths = 5.0 * w * h * spp
    value = paths / (ms_step * 1e-3) / 1e6
    # where the last timed step went on every rank (CUDA events on each rank's stream)
    keys = ("render", "exchange", "solve", "copy", "exchange+solve")
    mine = torch.tensor([R.last_ms.get(k, 0.0) for k in keys], dtype=torch.float64, device="cuda")
    per_rank = [torch.zeros_like(mine) for _ in range(world)] if world > 1 else [mine]
    if world > 1:
        dist.all_gather(per_rank, mine)
    per_rank_ms = {k: [round(float(t[i]), 2) for t in per_rank] for i, k in enumerate(keys) if any(float(t[i]) != 0.0 for t in per_rank)}

    # N > 1: the N-rank image against the same samples rendered by ONE GPU (same per-pixel RNG streams, so equal up to the
    # order of the floating-point accumulation) — the driver's scaling run thereby checks the sharded result, not just its speed
    mcheck = None
    if world > 1:
        cspp = 2
        R.render(cspp, precision=args.precision, offsets=args.offsets, seed=12345)
        barrier()
        if rank == 0:
            multi_out = [R.acc[0].clone(), R.out.clone()]
            one = g.gradient_path_render(sc, device=local, spp=cspp, precision=args.precision, offsets=args.offsets, seed=12345, outputs=("primal", "poisson"))
            mp = multi_out[0].cpu().numpy().reshape(h, w, 3); mo = multi_out[1].cpu().numpy().reshape(h, w, 3)
            rel = lambda a, b: float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300))
            mcheck = {"spp": cspp, "rel_l2_primal": rel(mp, one.primal), "rel_l2_poisson": rel(mo, one.poisson)}
            mcheck["ok"] = bool(mcheck["rel_l2_primal"] < 1e-9 and mcheck["rel_l2_poisson"] < 1e-7)
        barrier()

    # ---- end to end through the C ABI with host buffers (scene upload + BVH + render + 4 images back)
    def e2e_step():
        sc2 = g.Scene(sc.a)
        y0, y1 = R.bands[rank]
        if world == 1:
            r = g.gradient_path_render(sc2, device=local, spp=spp, precision=args.precision, offsets=args.offsets)
            sc2.close()
            return r
        R2 = multi.DistributedRenderer(sc2, local, dist_, torch, solver=args.solver, **share)
        r = R2.render(spp, precision=args.precision, to_host=True, offsets=args.offsets)
        sc2.close()
        return r
    scene_bytes = sum(v.nbytes for k, v in sc.a.items())
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        e2e_step()
    barrier()
    e2e_ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / args.steps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    e2e_val = paths / (float(e2e_ms.item()) * 1e-3) / 1e6

    # ---- the reference's own precision (Real = double, lajolla.h:23): the same workload in the FP64 mode that carries the
    # bit-level parity evidence, device-resident and end to end, a bounded number of full-size steps
    f64 = None
    if args.precision != "f64" and args.f64_steps > 0:
        R.render(max(1, spp // 64), precision="f64", offsets=args.offsets)       # warm-up: kernels + allocations of the double records
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for _ in range(args.f64_steps):
            R.render(spp, precision="f64", offsets=args.offsets)
        f1.record()
        barrier()
        fms = torch.tensor([f0.elapsed_time(f1) / args.f64_steps], dtype=torch.float64, device="cuda")
        saved = args.precision
        args.precision = "f64"
        t0 = time.perf_counter()
        e2e_step()
        barrier()
        fe = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device="cuda")
        args.precision = saved
        if world > 1:
            dist.all_reduce(fms, op=dist.ReduceOp.MAX); dist.all_reduce(fe, op=dist.ReduceOp.MAX)
        f64 = {"value": paths / (float(fms.item()) * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": float(fms.item()), "steps": args.f64_steps,
               "e2e": {"value": paths / (float(fe.item()) * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": float(fe.item()), "steps": 1},
               "note": "same workload in GDPT_PRECISION_F64 (double traversal tests + shading, -fmad=false): the mode whose per-sample records equal the compiled reference's"}

    # (every rank takes part: at N > 1 a render is a collective)
    # kernel durations: a separate pass of the same workload at fewer spp with the host-synchronised wave loop and CUDA events
    # around every k_trace / shade launch on the launching stream (in the timed region above two waves overlap on two
    # streams without read-backs, where per-launch events would time the neighbour's kernels too)
    check(lib.gdpt_set_option(R.ctx, b"time_kernels", 1))
    roof_spp = min(spp, 32)
    R.render(2, precision=args.precision, offsets=args.offsets)
    barrier()
    r0 = stats()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    R.render(roof_spp, precision=args.precision, offsets=args.offsets)
    g1.record()
    barrier()
    r1 = stats()
    check(lib.gdpt_set_option(R.ctx, b"time_kernels", 0))
    ms_pass = g0.elapsed_time(g1)
    rays = r1["rays_traced"] - r0["rays_traced"]
    ms_trace = r1["ms_trace"] - r0["ms_trace"]
    ms_shade = r1["ms_shade"] - r0["ms_shade"]
    rays_step = (s1["rays_traced"] - s0["rays_traced"]) / args.steps

    if rank == 0:
        # ---- roofline of the traversal kernel (k_trace): algorithmic bytes per ray from a counting pass
        peaks, how = measured_peaks()
        p = _params(sc, spp=1, precision=args.precision)
        cnt = np.zeros(3)
        check(lib.gdpt_traversal_counts(R.ctx, C.byref(p), _capi.as_dp(cnt)))
        bytes_per_ray = 32 + 20 + 128.0 * cnt[0] + 48.0 * cnt[1]
        achieved = rays * bytes_per_ray / (ms_trace * 1e-3) / 1e9 if ms_trace > 0 else None
        l2 = np.zeros(1)
        try:
            check(lib.gdpt_measure_l2_read(local, 48, 40, _capi.as_dp(l2)))     # 48 MB buffer, L2-resident
        except Exception:
            l2[0] = 0.0
        # committed ncu evidence for the same kernel (profiles/r2y_k_trace_bounce.json (final state of the kernel) and r2v_k_trace_primary.json; one --set full capture per launch type)
        ncu = {}
        for tag in ("bounce", "primary"):
            try:
                ncu[tag] = json.load(open(os.path.join(ROOT, "profiles", ("r2y_k_trace_%s.json" if tag == "bounce" else "r2v_k_trace_%s.json") % tag)))
            except Exception:
                pass
        nb = ncu.get("bounce")
        issue_peak = 148 * 4 * (clocks["sm_mhz"] or peaks.get("sm_max_mhz", 1965.0)) * 1e6     # warp instructions / s
        l2_peak = float(l2[0]) or None
        roof = {"bound": "l2", "kernel": "k_trace" if args.offsets == "reference" else "k_shift (ray casts + shading in one kernel; bytes = its rays x the traversal bytes per ray)",
                "achieved": achieved, "peak": l2_peak, "unit": "GB/s", "frac": (achieved / l2_peak) if achieved and l2_peak else None,
                "peak_source": "live L2 read microbenchmark in this run (gdpt_measure_l2_read: 48 MB buffer, ld.global.cg, all SMs); MEASURED_PEAKS.json has no L2 figure",
                "traffic": nb["dram_bytes"] if nb else None, "traffic_source": "committed ncu capture (profiles/), not this run",
                "bytes_per_ray": bytes_per_ray, "nodes_per_ray": cnt[0], "prims_per_ray": cnt[1],
                "hbm": {"peak": peaks["hbm_gbs"], "peak_source": how, "frac": achieved / peaks["hbm_gbs"] if achieved else None,
                        "note": "not a bound: the scene is cache-resident, DRAM sees only the ray / hit / permutation streams (`traffic`)"},
                "issue": ({"issue_active_frac": nb["issue_active_pct"] / 100.0, "threads_per_inst": nb["threads_per_inst"],
                           "lane_slot_frac": nb["issue_active_pct"] / 100.0 * nb["threads_per_inst"] / 32.0,
                           "issue_peak_winst_per_s": issue_peak, "source": "committed ncu capture of a bounce launch (profiles/), not this run"}
                          if nb else None),
                "note": "algorithmic bytes = 32 B ray + 20 B hit + node bytes x nodes + 48 B x primitives per ray (counting pass of this run), served by L1/L2: "
                        "the binding limits are L2/L1 bandwidth-latency and issue slots under divergence (DESIGN.md 4.1)",
                "kernel_timing_pass": {"spp": roof_spp, "ms": ms_pass, "ms_k_trace": ms_trace, "ms_shade_kernels": ms_shade, "rays": rays,
                                       "how": "host-synchronised wave loop, CUDA events around every launch on the launching stream"},
                "share_of_step": {"k_trace": ms_trace / ms_pass, "shade_kernels": ms_shade / ms_pass},
                "rays_per_step": rays_step,
                "mrays_per_s": rays / (ms_trace * 1e-3) / 1e6 if ms_trace > 0 else None}
        line = {"metric": metric_name(w, h), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": args.precision, "data": "synthetic",
                "config": {"workload": workload_name(args.scene, w, h, spp, args.offsets) + (", synthetic camera %d" % args.camera if args.camera else ""), "row_bands": world,
                           "precision": args.precision + " traversal/shading, f64 accumulation + Poisson",
                           "poisson_solver": R.solver if world > 1 else "single GPU",
                           "l2_flush": "inputs larger than L2: each step streams %.1f GB of path state" % (paths / 5 * 900 / 1e9)},
                "clocks": clocks, "gpu_launches": int(launches.item()),
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(scene_bytes), "d2h_bytes_per_step": int(4 * w * h * 3 * 8),
                        "ms_per_step": float(e2e_ms.item())},
                "roofline": roof,
                "msamples_per_s": value / 5.0}
        if f64:
            line["f64"] = f64
        line["per_rank_ms"] = per_rank_ms
        if mcheck:
            line["multi_gpu_check"] = mcheck
        line["config"]["film_sharing"] = ("interleaved 4-row strips" if getattr(R, "interleave", False) else "contiguous row bands") if world > 1 else "one GPU"
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args, args.scene, spp)
            line["equal_time"] = equal_time_block(g, local)
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_STDOUT_FD = None


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else printed by libraries during the run
    (NCCL's version banner, the reference's progress bar) was sent to stderr by main()."""
    data = (json.dumps(line) + "\n").encode()
    os.write(_STDOUT_FD if _STDOUT_FD is not None else 1, data)


def main():
    global _STDOUT_FD
    args = parse()
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_native(args)


if __name__ == "__main__":
    main()
