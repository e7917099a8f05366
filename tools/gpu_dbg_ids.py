import sys, numpy as np
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
name = sys.argv[1] if len(sys.argv) > 1 else "cbox_gdpt_64"
z = np.load("tests/golden/intersect_%s.npz" % name)
sc = g.Scene.named(name)
for key in ("", "2"):
    rays = z["rays" + key]; ids_w = z["ids" + key]; vtx_w = z["vtx" + key]
    ids, vtx = g.intersect(sc, rays, None if key else z["rdiff"], precision="f64")
    bad = np.where((ids != ids_w).any(axis=1))[0]
    print(name, key or "1", "mismatch", len(bad), "of", len(ids))
    for i in bad[:12]:
        print(" ray", rays[i], "\n   gpu", ids[i], "t", vtx[i, 22], "pos", vtx[i, :3], "\n   ref", ids_w[i], "t", vtx_w[i, 22], "pos", vtx_w[i, :3])
