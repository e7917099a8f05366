"""GPU tuning sweep: time the trace-kernel variants (GDPT_TRACE_VARIANT) on the bench scene."""
import os, sys, time
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
variants = [int(v) for v in sys.argv[1].split(",")] if len(sys.argv) > 1 else [0, 1, 2, 3, 10, 13, 100, 103, 110, 113, 200, 203]
scene = sys.argv[2] if len(sys.argv) > 2 else "sponza_1080p"
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 8
sc = g.Scene.named(scene)
sorts = [int(v) for v in sys.argv[4].split(",")] if len(sys.argv) > 4 else [0, 1]
ref = None
for v, srt in [(v, s_) for v in variants for s_ in sorts]:
    os.environ["GDPT_TRACE_VARIANT"] = str(v)
    os.environ["GDPT_SORT_RAYS"] = str(srt)
    sc.close()
    g.gradient_path_render(sc, precision="f32", spp=2)          # warm-up (also creates the context)
    best = None
    for _ in range(3):
        r = g.gradient_path_render(sc, precision="f32", spp=spp)
        st = r["stats"]
        if best is None or st["ms_trace"] < best["ms_trace"]: best = st
    m = float(r.primal.mean())
    if ref is None: ref = m
    print("variant %4d sort %d  trace %8.2f ms  shade %7.2f ms  Mrays/s %8.1f  primal mean %.6f %s" % (
        v, srt, best["ms_trace"], best["ms_shade"], best["rays_traced"] / best["ms_trace"] / 1e3, m, "" if abs(m - ref) < 1e-4 * abs(ref) else "MISMATCH"), flush=True)
