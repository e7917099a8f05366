"""GPU tuning sweep: GDPT_TAIL_PATHS (when the single-launch tail kernel takes over) on the bench scene."""
import os, sys
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
vals = [int(v) for v in sys.argv[1].split(",")] if len(sys.argv) > 1 else [0, 4096, 16384, 65536, 262144, 1048576]
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
sc = g.Scene.named("sponza_1080p")
for v in vals:
    os.environ["GDPT_TAIL_PATHS"] = str(v)
    sc.close()
    g.gradient_path_render(sc, precision="f32", spp=2)
    best = None
    for _ in range(3):
        r = g.gradient_path_render(sc, precision="f32", spp=spp)
        st = r["stats"]; acc = st["ms_total"] - st["ms_poisson"] - st["ms_copy"]
        if best is None or acc < best[0]: best = (acc, st)
    acc, st = best
    print("tail_paths %8d  accumulate %8.2f ms  trace %7.2f  shade %7.2f  launches %5d  Msamples/s %7.1f  mean %.6f" % (
        v, acc, st["ms_trace"], st["ms_shade"], st["kernel_launches"], 1920 * 1080 * spp / acc / 1e3, float(r.primal.mean())), flush=True)
