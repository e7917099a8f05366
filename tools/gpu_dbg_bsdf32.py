import sys, numpy as np
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
name = sys.argv[1] if len(sys.argv) > 1 else "matpreview_small"
sc = g.Scene.named(name)
rng = np.random.default_rng(3)
def unit(v): return v / np.linalg.norm(v, axis=-1, keepdims=True)
n = 4000
nrm = unit(rng.standard_normal((n, 3))); t = unit(np.cross(nrm, unit(rng.standard_normal((n, 3))))); b = np.cross(nrm, t)
gn = nrm.copy()
uv = rng.random((n, 2)); fp = np.zeros((n, 1))
vtx = np.concatenate([gn, t, b, nrm, uv, fp], axis=1)
din = unit(rng.standard_normal((n, 3))); dout = unit(rng.standard_normal((n, 3)))
dirs = np.concatenate([din, dout], axis=1); rnd = rng.random((n, 3))
print("materials", sc.a["materials"][:, 0], "eta", sc.a["materials"][:, 1])
for m in range(sc.a["materials"].shape[0]):
    o64 = g.bsdf(sc, m, vtx, dirs, rnd, precision="f64"); o32 = g.bsdf(sc, m, vtx, dirs, rnd, precision="f32")
    nan32 = ~np.isfinite(o32); nan64 = ~np.isfinite(o64)
    print("mat", m, "type", int(sc.a["materials"][m, 0]), "f64 non-finite per column", nan64.sum(0), "f32", nan32.sum(0))
    bad = np.nonzero(nan32.any(1) & ~nan64.any(1))[0][:3]
    for i in bad: print("   ", i, "f32", o32[i], "f64", o64[i], "cos_in", (din[i] * nrm[i]).sum(), "cos_out", (dout[i] * nrm[i]).sum())
