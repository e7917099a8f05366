"""relMSE at equal time (the second half of BASELINE.json's metric): GPU gradpath (primal and Poisson-reconstructed), GPU
path (NEE + MIS) and the reference's CPU gradpath / path on the box's host cores, each given the same wall-clock budget on
the shipped Cornell box (scenes/cbox/cbox_gdpt.xml, 512x512).  relMSE = mean((I - R)^2 / (R^2 + 1e-3)).

Every estimator is compared with ITS OWN converged image, rendered here on the GPU in f64-accumulated f32 mode at REF_SPP:
the two reference integrators do not converge to the same picture (as shipped, `path` adds BSDF-sampled light hits
without the MIS weight w2, path_tracing.h:228-232, and gradpath has no next-event estimation), so a common target would
measure their bias, not their noise.  Usage: python tools/gpu_equal_time.py [scene] [ref_spp] [shift]
With a third argument only the shift-mapped gradpath (offsets="shift") is measured: primal and Poisson image
against the converged primal of that mode (its Poisson image converges to the same picture)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gradient_based_path_tracing_b200 as g

name = sys.argv[1] if len(sys.argv) > 1 else "cbox_gdpt"
REF_SPP = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
BUDGETS = [0.5, 2.0, 8.0]


def relmse(i, r):
    return float(np.mean((i - r) ** 2 / (r ** 2 + 1e-3)))


sc = g.Scene.named(name)
w, h = sc.width, sc.height
if len(sys.argv) > 3:
    g.gradient_path_render(sc, spp=1, precision="f32", offsets="shift")
    t = time.perf_counter(); Rs = g.gradient_path_render(sc, spp=REF_SPP, precision="f32", seed=99, offsets="shift"); ts = time.perf_counter() - t
    rate_s = REF_SPP / ts
    print("%s %dx%d shift-mapped gradpath: reference at %d spp in %.1f s (%.1f Msamples/s); converged Poisson vs primal relMSE %.2e"
          % (name, w, h, REF_SPP, ts, w * h * rate_s / 1e6, relmse(Rs.poisson, Rs.primal)), flush=True)
    for T in BUDGETS:
        n = max(1, int(rate_s * T))
        t = time.perf_counter(); a = g.gradient_path_render(sc, spp=n, precision="f32", seed=7, offsets="shift"); ta = time.perf_counter() - t
        print("%7.2fs | %5d spp %.2fs | primal %.2e | Poisson %.2e | ratio %.2f" % (T, n, ta, relmse(a.primal, Rs.primal), relmse(a.poisson, Rs.primal),
              relmse(a.primal, Rs.primal) / relmse(a.poisson, Rs.primal)), flush=True)
    sys.exit(0)
g.gradient_path_render(sc, spp=1, precision="f32")
g.path_render(sc, spp=1, precision="f32")
t = time.perf_counter(); Rg = g.gradient_path_render(sc, spp=REF_SPP, precision="f32", seed=99); tg = time.perf_counter() - t
t = time.perf_counter(); Rp = g.path_render(sc, spp=REF_SPP, precision="f32", seed=99); tp = time.perf_counter() - t
rate_g, rate_p = REF_SPP / tg, REF_SPP / tp       # spp per second at this film size
print("%s %dx%d: references at %d spp: gradpath %.1f s (%.1f Msamples/s = %.1f Mpaths/s), path %.1f s (%.1f Msamples/s)"
      % (name, w, h, REF_SPP, tg, w * h * rate_g / 1e6, 5 * w * h * rate_g / 1e6, tp, w * h * rate_p / 1e6), flush=True)
st = g.stats(sc)
print("path stats at %d spp: trace %.1f ms, shade %.1f ms, total %.1f ms, %d launches, %.2f rays per sample" % (REF_SPP, st["ms_trace"], st["ms_shade"], st["ms_total"], st["kernel_launches"], st["rays_traced"] / (w * h * REF_SPP)))
print("reference means: gradpath primal %.5f, gradpath poisson %.5f, path %.5f" % (Rg.primal.mean(), Rg.poisson.mean(), Rp.mean()))

cpu = None
try:
    from oracle import ref
    if ref.available() and ref.scenes_available():
        cpu = ref.RefScene(ref.scene_path(name))
        cores = os.cpu_count()
        cpu.set_options(spp=1)
        t = time.perf_counter(); cpu.render(); cg = time.perf_counter() - t          # seconds per spp, gradpath, all host threads
        t = time.perf_counter(); cpu.render_path(); cp = time.perf_counter() - t
        print("CPU reference on %d host threads: gradpath %.2f s/spp, path %.2f s/spp" % (cores, cg, cp), flush=True)
except Exception as e:  # the oracle is optional here
    print("CPU reference unavailable:", e)

print("%8s | %22s | %22s | %22s | %22s | %22s" % ("budget", "GPU gradpath primal", "GPU gradpath Poisson", "GPU path", "CPU gradpath Poisson", "CPU path"))
for T in BUDGETS:
    row = []
    sg = max(1, int(rate_g * T)); sp = max(1, int(rate_p * T))
    t = time.perf_counter(); a = g.gradient_path_render(sc, spp=sg, precision="f32", seed=7); ta = time.perf_counter() - t
    row.append("%5d spp %.2fs %.2e" % (sg, ta, relmse(a.primal, Rg.primal)))
    row.append("%5d spp %.2fs %.2e" % (sg, ta, relmse(a.poisson, Rg.poisson)))
    t = time.perf_counter(); b = g.path_render(sc, spp=sp, precision="f32", seed=7); tb = time.perf_counter() - t
    row.append("%5d spp %.2fs %.2e" % (sp, tb, relmse(b, Rp)))
    if cpu is not None:
        s1 = max(1, int(round(T / cg))); s2 = max(1, int(round(T / cp)))
        cpu.set_options(spp=s1); t = time.perf_counter(); c = cpu.render(); tc = time.perf_counter() - t
        row.append("%5d spp %.2fs %.2e" % (s1, tc, relmse(c["poisson"], Rg.poisson)))
        cpu.set_options(spp=s2); t = time.perf_counter(); d = cpu.render_path(); td = time.perf_counter() - t
        row.append("%5d spp %.2fs %.2e" % (s2, td, relmse(d, Rp)))
    else:
        row += ["n/a", "n/a"]
    print("%7.2fs | " % T + " | ".join("%22s" % r for r in row), flush=True)
