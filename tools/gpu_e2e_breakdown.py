"""Where the end-to-end step spends its time beyond the device-resident render (Sponza 1080p, low spp so overheads show)."""
import sys, time
sys.path.insert(0, '.')
import numpy as np
import gradient_based_path_tracing_b200 as g
sc = g.Scene.named("sponza_1080p")
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
for k in range(4):
    t0 = time.perf_counter()
    sc2 = g.Scene(sc.a)
    t1 = time.perf_counter()
    ctx = sc2.context(0)
    t2 = time.perf_counter()
    r = g.gradient_path_render(sc2, device=0, spp=spp, precision="f32")
    t3 = time.perf_counter()
    sc2.close()
    t4 = time.perf_counter()
    print("step %d: Scene() %.1f  create %.1f  render %.1f (stats: total %.1f trace %.1f shade %.1f poisson %.1f copy %.1f)  close %.1f  sum %.1f ms" % (
        k, 1e3 * (t1 - t0), 1e3 * (t2 - t1), 1e3 * (t3 - t2), r.stats["ms_total"], r.stats["ms_trace"], r.stats["ms_shade"], r.stats["ms_poisson"], r.stats["ms_copy"],
        1e3 * (t4 - t3), 1e3 * (t4 - t0)))
