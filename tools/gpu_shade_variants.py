"""GPU tuning sweep: launch bounds of the diffuse-only float shade kernels (GDPT_SHADE_VARIANT = 10*primary + bounce)."""
import os, sys
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
vals = [int(v) for v in sys.argv[1].split(",")] if len(sys.argv) > 1 else [0, 1, 2, 3, 10, 20, 30, 22]
sc = g.Scene.named("sponza_1080p")
for v in vals:
    os.environ["GDPT_SHADE_VARIANT"] = str(v)
    sc.close()
    g.gradient_path_render(sc, precision="f32", spp=2)
    best = None
    for _ in range(3):
        r = g.gradient_path_render(sc, precision="f32", spp=8)
        st = r["stats"]
        if best is None or st["ms_shade"] < best["ms_shade"]: best = st
    print("shade_variant %3d  shade %7.2f ms  trace %7.2f ms  mean %.6f" % (v, best["ms_shade"], best["ms_trace"], float(r.primal.mean())), flush=True)
