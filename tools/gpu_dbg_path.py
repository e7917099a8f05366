import sys, numpy as np
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
name = sys.argv[1] if len(sys.argv) > 1 else "cbox_gdpt_d2"
z = np.load("tests/golden/path_%s.npz" % name)
sc = g.Scene.named(name)
got = g.path_samples(sc, z["xs"], z["ys"], z["streams"], precision="f64")
want = z["rad"]
bad = (np.abs(got - want) > 1e-9 * (1 + np.abs(want))).any(axis=1)
print(name, "bad", bad.sum(), "of", bad.size, "mean got", got.mean(), "want", want.mean(), "max_depth", z["max_depth"], "rr", z["rr_depth"])
idx = np.nonzero(bad)[0][:12]
for i in idx:
    print(i, z["xs"][i], z["ys"][i], "got", got[i], "want", want[i], "ratio", got[i] / np.where(want[i] != 0, want[i], 1))
for md in (1, 2, 3):
    got = g.path_samples(sc, z["xs"][:8], z["ys"][:8], z["streams"][:8], precision="f64", max_depth=md)
    print("max_depth", md, got[:8, 0])
