import sys, numpy as np
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
from oracle import ref
np.set_printoptions(linewidth=200, precision=6)
sc = g.Scene.named("cbox_gdpt_64")
rs = ref.RefScene(ref.scene_path("cbox_gdpt_64"), threads=2)
rng = np.random.default_rng(0)
xs, ys = rng.integers(0, 64, 256), rng.integers(0, 64, 256)
st = np.arange(256, dtype=np.uint64)
rec = g.grad_samples(sc, xs, ys, st, precision="f64")
want = np.array([rs.grad_sample(int(x), int(y), int(s)) for x, y, s in zip(xs, ys, st)])
want2 = np.array([rs.grad_sample(int(x), int(y), int(s)) for x, y, s in zip(xs, ys, st)])
print("reference repeatable:", np.array_equal(np.nan_to_num(want), np.nan_to_num(want2)), "nan rows", np.isnan(want).any(axis=1).sum())
bad = (np.abs(rec - want) > 1e-9 * (1 + np.abs(want))).any(axis=1)
print("bad", bad.sum(), np.nonzero(bad)[0])
for i in np.nonzero(bad)[0]:
    print(i, xs[i], ys[i]); print("  got ", rec[i]); print("  want", want[i])
z = np.load("tests/golden/records_cbox_gdpt_64.npz")
rec2 = g.grad_samples(sc, z["xs"], z["ys"], z["streams"], precision="f64")
w = z["rec"]; b2 = (np.abs(rec2 - w) > 1e-9 * (1 + np.abs(w))).any(axis=1)
print("golden fixture: bad", b2.sum(), "of", len(w))
w3 = np.array([rs.grad_sample(int(x), int(y), int(s)) for x, y, s in zip(z["xs"][:300], z["ys"][:300], z["streams"][:300])])
print("reference now vs fixture (first 300): equal", np.array_equal(np.nan_to_num(w3), np.nan_to_num(w[:300])))
