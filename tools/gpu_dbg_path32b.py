import sys, numpy as np
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
name = "matpreview_small"
z = np.load("tests/golden/path_%s.npz" % name)
sc = g.Scene.named(name)
idx = np.array([2, 9, 14, 20, 25, 26, 0, 1, 3])
xs, ys, st = z["xs"][idx], z["ys"][idx], z["streams"][idx]
for md in (1, 2, 3, 4, -1):
    a = g.path_samples(sc, xs, ys, st, precision="f32", max_depth=md)
    b = g.path_samples(sc, xs, ys, st, precision="f64", max_depth=md)
    print("max_depth", md, "f32", a[:, 0], "\n              f64", b[:, 0])
# primary hits: material of each sample's pixel centre
sp = np.stack([(xs + 0.5) / sc.width, (ys + 0.5) / sc.height], axis=1)
rays6 = g.sample_primary(sc, sp)
rays = np.concatenate([rays6, np.zeros((len(idx), 1)), np.full((len(idx), 1), np.inf)], axis=1)
ids, vtx = g.intersect(sc, rays)
print("primary ids (shape, prim, material):", ids.tolist())
ids32, vtx32 = g.intersect(sc, rays, precision="f32")
print("f32 vertex finite:", np.isfinite(vtx32).all(axis=1), "uv", vtx32[:, 17:19].tolist())
print("textures", sc.a["textures"][:, :2].tolist())
