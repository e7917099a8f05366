import sys, numpy as np
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
name = sys.argv[1] if len(sys.argv) > 1 else "matpreview_small"
z = np.load("tests/golden/path_%s.npz" % name)
sc = g.Scene.named(name)
want = z["rad"]
for prec in ("f64", "f32"):
    got = g.path_samples(sc, z["xs"], z["ys"], z["streams"], precision=prec)
    fin = np.isfinite(got).all(axis=1)
    close = (np.abs(got - want) <= 2e-3 * (1 + np.abs(want))).all(axis=1)
    print(name, prec, "non-finite", (~fin).sum(), "close", close.mean(), "mean got", got[fin].mean(), "want", want.mean())
    bad = np.nonzero(~close)[0][:6]
    for i in bad: print("   ", i, got[i], want[i])
    img = g.path_render(sc, spp=4, precision=prec)
    print("   image: non-finite pixels", (~np.isfinite(img)).any(axis=2).sum(), "mean", np.nanmean(img))
