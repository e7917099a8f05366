"""GPU tuning sweep: resolution of the ray-sort key (GDPT_SORT_BITS = "cell bits per axis,direction bits per axis")."""
import os, sys
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
combos = sys.argv[1].split(";") if len(sys.argv) > 1 else ["off", "4,2", "5,0", "5,2", "5,3", "5,4", "6,2", "6,3", "4,4", "7,2"]
sc = g.Scene.named("sponza_1080p")
for cb in combos:
    if cb == "off":
        os.environ["GDPT_SORT_RAYS"] = "0"
    else:
        os.environ["GDPT_SORT_RAYS"] = "1"; os.environ["GDPT_SORT_BITS"] = cb
    sc.close()
    g.gradient_path_render(sc, precision="f32", spp=2)
    best = None
    for _ in range(3):
        r = g.gradient_path_render(sc, precision="f32", spp=8)
        st = r["stats"]; acc = st["ms_total"] - st["ms_poisson"] - st["ms_copy"]
        if best is None or acc < best[0]: best = (acc, st)
    acc, st = best
    print("sort %-5s accumulate %7.2f ms  trace(+sort) %7.2f  shade %6.2f  mean %.6f" % (cb, acc, st["ms_trace"], st["ms_shade"], float(r.primal.mean())), flush=True)
