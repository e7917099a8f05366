import sys, numpy as np
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
np.set_printoptions(linewidth=200, precision=6)
for name in sys.argv[1:] or ["cbox_gdpt_64"]:
    z = np.load("tests/golden/records_%s.npz" % name)
    sc = g.Scene.named(name)
    rec = g.grad_samples(sc, z["xs"], z["ys"], z["streams"], precision="f64")
    want = z["rec"]
    fin = np.isfinite(want).all(axis=1)
    err = np.abs(rec - want) / (1e-12 + np.abs(want))
    bad = np.where(fin & (err > 1e-7).any(axis=1))[0]
    exact = (rec == want).all(axis=1).sum()
    print(name, "bad", len(bad), "of", len(rec), "bit-identical records", exact, "median err", np.median(err[fin]))
    for i in bad[:6]:
        print(" i", i, "xy", z["xs"][i], z["ys"][i], "\n   gpu", rec[i], "\n   ref", want[i])
