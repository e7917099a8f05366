"""GPU probe: where the end-to-end time of one reference-facing call goes (GDPT_VERBOSE prints gdpt_create's phases)."""
import os, sys, time
os.environ["GDPT_VERBOSE"] = "1"
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
base = g.Scene.named("sponza_1080p")
for rep in range(2):
    t0 = time.perf_counter(); sc = g.Scene(base.a); t1 = time.perf_counter()
    sc.context(0); t2 = time.perf_counter()
    r = g.gradient_path_render(sc, precision="f32", spp=8); t3 = time.perf_counter()
    r = g.gradient_path_render(sc, precision="f32", spp=8); t4 = time.perf_counter()
    sc.close(); t5 = time.perf_counter()
    print("rep %d: Scene() %.3f  gdpt_create %.3f  first render(8spp) %.3f  second render %.3f  destroy %.3f   stats %s" % (
        rep, t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, {k: round(v, 1) for k, v in r["stats"].items() if k.startswith("ms_")}), flush=True)
