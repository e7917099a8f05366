"""B200 measurements for the smallgdpt entry point: closeness to the reference program's fixtures, and timings of both stream
layouts next to the reference program on this box's host cores.  Output -> gpurun_out/r2_smallgdpt.txt"""
import glob
import os
import sys
import time

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
from gradient_based_path_tracing_b200 import smallgdpt  # noqa: E402
from oracle import smallgdpt as ref  # noqa: E402

os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
out = open(os.path.join(ROOT, "gpurun_out", "r2_smallgdpt.txt"), "w")


def say(*a):
    s = " ".join(str(v) for v in a)
    print(s); out.write(s + "\n"); out.flush()


say("# gdpt_smallgdpt_render on B200 against the reference program small_gdpt.cpp (fixtures: its PFM files)")
for path in sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "smallgdpt_*.npz"))):
    gold = np.load(path)
    r = smallgdpt.render(int(gold["width"]), int(gold["height"]), int(gold["spp"]), rng="rows")
    mine = smallgdpt.pfm_images(r)
    for n in ref.NAMES:
        a, b = mine[n], gold[n]
        say(os.path.basename(path), n, "identical float32 values: %.6f" % float((a == b).mean()),
            "max |d|/(|ref|+1e-3): %.3g" % float((np.abs(a.astype(np.float64) - b) / (np.abs(b) + 1e-3)).max()))
w, h = 1024, 768
for spp in (4, 16):
    smallgdpt.render(w, h, 1, rng="rows")
    t0 = time.perf_counter(); a = smallgdpt.render(w, h, spp, rng="rows"); ta = time.perf_counter() - t0
    t0 = time.perf_counter(); b = smallgdpt.render(w, h, spp, rng="pixels"); tb = time.perf_counter() - t0
    say("film %dx%d, %d spp: rows layout kernel %.1f ms (call %.1f ms), pixels layout kernel %.1f ms (call %.1f ms)" %
        (w, h, spp, a["render_ms"], ta * 1e3, b["render_ms"], tb * 1e3),
        "-> %.1f / %.1f Msamples/s device-timed" % (w * h * spp / a["render_ms"] / 1e3, w * h * spp / b["render_ms"] / 1e3))
    if ref.available():
        t = ref.run(w, h, spp)
        say("  reference program, %d host threads: %.2f s whole program -> %.2f Msamples/s" % (os.cpu_count(), t["seconds"], w * h * spp / t["seconds"] / 1e6))
        m = smallgdpt.pfm_images(a)
        keep = ~ref.stale_read_mask(w, h)
        for n in ref.NAMES:
            x, y = m[n], t[n]
            if n in ("image_dx", "image_dy"):
                say("   ", n, "outside the stale-read mask: identical %.6f, worst %.3g; whole film: identical %.6f" %
                    (float((x[keep] == y[keep]).mean()), float((np.abs(x[keep].astype(np.float64) - y[keep]) / (np.abs(y[keep]) + 1e-3)).max()), float((x == y).mean())))
            else:
                say("   ", n, "identical %.6f, worst %.3g" % (float((x == y).mean()), float((np.abs(x.astype(np.float64) - y) / (np.abs(y) + 1e-3)).max())))
b = smallgdpt.render(w, h, 256, rng="pixels")
say("film %dx%d, 256 spp, pixels layout: kernel %.1f ms -> %.1f Msamples/s" % (w, h, b["render_ms"], w * h * 256 / b["render_ms"] / 1e3))
