import sys, time, numpy as np
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
from oracle import ref
print("devices", g._capi.lib().gdpt_device_count())
out, raw = g.pcg_fill(0, 4); print("pcg", out, raw)
sc = g.Scene.named("cbox_gdpt")
t=time.time(); sc.context(0); print("create", time.time()-t)
print("bounds", g.scene_bounds(sc))
rs = ref.RefScene(ref.scene_path("cbox_gdpt"))
rng = np.random.default_rng(0)
n = 4000
xs = rng.integers(0, 512, n); ys = rng.integers(0, 512, n); st = np.arange(n, dtype=np.uint64) + 7
rec = g.grad_samples(sc, xs, ys, st, precision="f64")
refrec = np.array([rs.grad_sample(int(x), int(y), int(s)) for x, y, s in zip(xs, ys, st)])
d = np.abs(rec - refrec) / (1e-12 + np.abs(refrec))
bad = (d.max(axis=1) > 1e-6)
print("f64 records: mismatching samples", bad.sum(), "of", n, "max rel", d.max())
for i in np.where(bad)[0][:5]: print(i, xs[i], ys[i], "\n gpu", rec[i], "\n ref", refrec[i])
rec32 = g.grad_samples(sc, xs, ys, st, precision="f32")
d32 = np.abs(rec32 - refrec) / (1e-6 + np.abs(refrec))
print("f32 records: >1e-3 mismatches", (d32.max(axis=1) > 1e-3).sum())
for prec in ("f64", "f32"):
    t=time.time(); r = g.gradient_path_render(sc, precision=prec, spp=4); dt=time.time()-t
    print(prec, "render 4spp", dt, {k: float(v.mean()) for k, v in r.items() if k != "stats"}, r["stats"])
t=time.time(); rr = rs.render(); print("ref render", time.time()-t, {k: float(v.mean()) for k, v in rr.items()})
