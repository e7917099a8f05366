"""GPU debug: pixel-seeded cbox64 image vs golden; run-to-run determinism; where the gradient differs."""
import sys, numpy as np
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
from oracle import cport
z = np.load("tests/golden/render_cbox64.npz")
sc = g.Scene.named("cbox_gdpt_64")
cx, cy = cport.merge(z["cx0"], z["cx1"], z["cy0"], z["cy1"])
rs = [g.gradient_path_render(sc, spp=2, precision="f64") for _ in range(3)]
for i, r in enumerate(rs):
    print("run", i, "primal==golden", np.allclose(r.primal, z["c"], rtol=1e-9, atol=1e-12),
          "dx close", np.isclose(r.dx, cx, rtol=1e-7, atol=1e-9).mean(), "dy close", np.isclose(r.dy, cy, rtol=1e-7, atol=1e-9).mean(),
          "same as run0 dx", np.isclose(r.dx, rs[0].dx, rtol=1e-12, atol=1e-14).mean())
r = rs[0]
bad = ~np.isclose(r.dx, cx, rtol=1e-7, atol=1e-9).all(axis=2)
ys, xs = np.nonzero(bad)
print("bad dx pixels", len(ys))
for y, x in list(zip(ys, xs))[:40]:
    print(y, x, r.dx[y, x], cx[y, x])
bad = ~np.isclose(r.dy, cy, rtol=1e-7, atol=1e-9).all(axis=2)
ys, xs = np.nonzero(bad)
print("bad dy pixels", len(ys))
for y, x in list(zip(ys, xs))[:40]:
    print(y, x, r.dy[y, x], cy[y, x])
