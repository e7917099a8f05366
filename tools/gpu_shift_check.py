"""Shift-mapped gradpath (offsets="shift") on the GPU: unbiasedness of the gradients against finite differences of the
converged primal, primal against the reference-mode primal, error of the Poisson image against the primal at equal spp,
throughput on Sponza."""
import sys, time
sys.path.insert(0, '.')
import numpy as np
import gradient_based_path_tracing_b200 as g

def relmse(a, ref): return float(np.mean((a - ref) ** 2 / (ref ** 2 + 1e-3)))
def rel_l2(a, b): return float(np.linalg.norm(a - b) / np.linalg.norm(b))

name = sys.argv[1] if len(sys.argv) > 1 else "cbox_gdpt_64"
prec = sys.argv[2] if len(sys.argv) > 2 else "f32"
sc = g.Scene.named(name)
big = int(sys.argv[4]) if len(sys.argv) > 4 else 16384
T = g.gradient_path_render(sc, spp=big, precision=prec, offsets="shift", sample0=100000)
R = g.gradient_path_render(sc, spp=big, precision=prec, sample0=200000)
print("%s %s: converged primal mean shift %.5f reference %.5f, rel-L2 between them %.4f" % (name, prec, T.primal.mean(), R.primal.mean(), rel_l2(T.primal, R.primal)))
Dx = np.zeros_like(T.primal); Dx[:, 1:] = T.primal[:, 1:] - T.primal[:, :-1]
Dy = np.zeros_like(T.primal); Dy[1:, :] = T.primal[1:, :] - T.primal[:-1, :]
print("gradients at %d spp vs finite differences of the converged primal (interior): shift dx %.4f dy %.4f | reference dx %.4f dy %.4f" % (
    big, rel_l2(T.dx[:, 1:], Dx[:, 1:]), rel_l2(T.dy[1:], Dy[1:]), rel_l2(R.dx[:, 1:], Dx[:, 1:]), rel_l2(R.dy[1:], Dy[1:])))
print("   column sums: dx %.5f vs %.5f ; dy %.5f vs %.5f" % (T.dx[:, 1:].sum(), Dx[:, 1:].sum(), T.dy[1:].sum(), Dy[1:].sum()))
print("   converged Poisson image vs converged primal: rel-L2 %.4f" % rel_l2(T.poisson, T.primal))
for n in (big // 64, big // 16, big // 4):
    a = g.gradient_path_render(sc, spp=n, precision=prec, offsets="shift", sample0=300000)
    print("   %7d spp: Poisson vs own primal rel-L2 %.4f, Poisson vs converged primal %.4f, primal vs converged primal %.4f" % (n, rel_l2(a.poisson, a.primal), rel_l2(a.poisson, T.primal), rel_l2(a.primal, T.primal)))
for spp in (16, 64, 256):
    a = g.gradient_path_render(sc, spp=spp, precision=prec, offsets="shift")
    b = g.gradient_path_render(sc, spp=spp, precision=prec)
    print("spp %4d: shift primal relMSE %.4e Poisson %.4e | reference-mode primal %.4e Poisson %.4e (each against the 16k primal of its mode)" % (
        spp, relmse(a.primal, T.primal), relmse(a.poisson, T.primal), relmse(b.primal, R.primal), relmse(b.poisson, R.primal)))
if len(sys.argv) > 3 and sys.argv[3] != "-":
    big_sc = g.Scene.named(sys.argv[3])
    for mode in ("shift", "reference"):
        g.gradient_path_render(big_sc, spp=2, precision="f32", offsets=mode)
        t = time.perf_counter(); r = g.gradient_path_render(big_sc, spp=8, precision="f32", offsets=mode); dt = time.perf_counter() - t
        s = r.stats
        print("%s %s: 8 spp device %.1f ms (trace %.1f shade %.1f) rays %.1f M -> %.1f Msamples/s, %.0f Mrays/s; primal mean %.5f" % (
            sys.argv[3], mode, s["ms_total"], s["ms_trace"], s["ms_shade"], s["rays_traced"] / 1e6,
            big_sc.width * big_sc.height * 8 / (s["ms_trace"] + s["ms_shade"]) / 1e3, s["rays_traced"] / (s["ms_trace"] + s["ms_shade"]) / 1e3, r.primal.mean()))
