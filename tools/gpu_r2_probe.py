"""GPU probe (round 2): wave scheduling A/B (host-synchronised vs device-side counts, 1 or 2 streams), samples per wave,
trace-kernel loop variants, f32 record error quantiles per scene, and the phases of one end-to-end call."""
import os, sys, time
sys.path.insert(0, '.')
import numpy as np
import gradient_based_path_tracing_b200 as g

what = sys.argv[1:] or ["sched", "records", "e2e"]
if "sched" in what:
    base = g.Scene.named("sponza_1080p")
    spp = 32
    def run(tag, opts, wave=0, prec="f32"):
        sc = g.Scene(base.a)
        for k, v in opts.items():
            sc.set_option(k, v)
        g.gradient_path_render(sc, spp=4, precision=prec, outputs=("primal",), wave_samples=wave)
        t0 = time.perf_counter()
        r = g.gradient_path_render(sc, spp=spp, precision=prec, outputs=("primal",), wave_samples=wave)
        dt = time.perf_counter() - t0
        print("%-44s %7.1f ms  %7.1f Mpaths/s   mean %.6f  launches %d" % (tag, dt * 1e3, 5 * 1920 * 1080 * spp / dt / 1e6, r.primal.mean(), r.stats["kernel_launches"]), flush=True)
        sc.close()
    run("default", {})
    for rep in range(3):
        run("f64, default (packed triangle records)", {}, prec="f64")
        run("f64, tri_records = 0 (index + vertex pools)", {"tri_records": 0}, prec="f64")
if "records" in what:
    for name in ["cbox_gdpt_64", "cbox_gdpt_d2", "veach_mi_small", "sponza_small", "disney_small"]:
        z = np.load("tests/golden/records_%s.npz" % name)
        sc = g.Scene.named(name)
        rec = g.grad_samples(sc, z["xs"], z["ys"], z["streams"], precision="f32")
        want = z["rec"]
        fin = np.isfinite(want).all(axis=1) & np.isfinite(rec).all(axis=1)
        err = np.abs(rec[fin] - want[fin]) / (1e-3 + np.abs(want[fin]))
        print("%-16s finite %.4f  frac(any field > 2e-3) %.4f  median %.2e  q90 %.2e q99 %.2e  radiance sums %s vs %s" % (
            name, fin.mean(), (err > 2e-3).any(axis=1).mean(), np.median(err), np.quantile(err, 0.9), np.quantile(err, 0.99),
            rec[fin][:, :3].sum(axis=0), want[fin][:, :3].sum(axis=0)), flush=True)
        sc.close()
if "e2e" in what:
    os.environ["GDPT_VERBOSE"] = "1"
    base = g.Scene.named("sponza_1080p")
    for rep in range(3):
        t0 = time.perf_counter(); sc = g.Scene(base.a); t1 = time.perf_counter()
        sc.context(0); t2 = time.perf_counter()
        r = g.gradient_path_render(sc, precision="f32", spp=8); t3 = time.perf_counter()
        r = g.gradient_path_render(sc, precision="f32", spp=8); t4 = time.perf_counter()
        sc.close(); t5 = time.perf_counter()
        print("rep %d: Scene() %.3f  gdpt_create %.3f  first render(8spp) %.3f  second render %.3f  destroy %.3f   stats %s" % (
            rep, t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, {k: round(v, 1) for k, v in r.stats.items() if k.startswith("ms_")}), flush=True)
if "bvh" in what:
    import ctypes as C
    from gradient_based_path_tracing_b200 import _capi
    from gradient_based_path_tracing_b200.render import _params
    base = g.Scene.named("sponza_1080p")
    os.environ["GDPT_VERBOSE"] = "1"
    for mode, split, radius, leaf, noself in (("host", "2", 8, 4, 1), ("host", "2", 8, 4, 1), ("host", "3", 8, 4, 1), ("host", "4", 8, 4, 1), ("host", "6", 8, 4, 1), ("host", "2", 8, 4, 1), ("host", "3", 8, 4, 1),
                                              ("host", "4", 8, 4, 1), ("gpu", "0", 8, 4, 1)):
        os.environ["GDPT_BVH"] = mode; os.environ["GDPT_BVH_SPLIT"] = split; os.environ["GDPT_BVH_SPLIT_DEVICE"] = split if mode == "gpu" else "0"
        os.environ["GDPT_PLOC_RADIUS"] = str(radius); os.environ["GDPT_BVH_MAX_LEAF"] = str(leaf); os.environ["GDPT_PLOC_NOSELF"] = str(noself)
        for rep in range(2):
            sc = g.Scene(base.a)
            t0 = time.perf_counter(); sc.context(0); t1 = time.perf_counter()
        nn, npr, depth, _, _ = sc.bvh_info(0)
        g.gradient_path_render(sc, spp=4, precision="f32", outputs=("primal",))
        t2 = time.perf_counter()
        r = g.gradient_path_render(sc, spp=32, precision="f32", outputs=("primal",))
        dt = time.perf_counter() - t2
        cnt = np.zeros(3)
        p = _params(sc, spp=1, precision="f32")
        _capi.check(_capi.lib().gdpt_traversal_counts(sc.context(0), C.byref(p), _capi.as_dp(cnt)))
        print("BVH %-4s split %s radius %d leaf %d noself %d: create %.1f ms, %d nodes, %d prim records, depth %d; 32 spp %.1f ms (%.1f Mpaths/s), %.2f nodes + %.2f prims per ray, mean %.6f" % (
            mode, split, radius, leaf, noself, (t1 - t0) * 1e3, nn, npr, depth, dt * 1e3, 5 * 1920 * 1080 * 32 / dt / 1e6, cnt[0], cnt[1], r.primal.mean()), flush=True)
        sc.close()
    del os.environ["GDPT_BVH"], os.environ["GDPT_BVH_SPLIT"], os.environ["GDPT_VERBOSE"], os.environ["GDPT_PLOC_RADIUS"], os.environ["GDPT_BVH_MAX_LEAF"], os.environ["GDPT_PLOC_NOSELF"], os.environ["GDPT_BVH_SPLIT_DEVICE"]
