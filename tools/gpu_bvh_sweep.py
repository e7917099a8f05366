"""GPU tuning sweep: BVH builder knobs (SAH primitive cost, maximum leaf size) on the bench scene."""
import os, sys, ctypes as C
sys.path.insert(0, '.')
import numpy as np
import gradient_based_path_tracing_b200 as g
from gradient_based_path_tracing_b200 import _capi
from gradient_based_path_tracing_b200.render import _params
combos = [tuple(c.split(":")) for c in sys.argv[1].split(",")] if len(sys.argv) > 1 else [("1", "4", "0"), ("1", "4", "0.25"), ("1", "4", "0.5"), ("1", "4", "1"), ("1", "4", "2"), ("1", "4", "4")]
sc = g.Scene.named("sponza_1080p")
for ci, ml, sp in combos:
    os.environ["GDPT_BVH_CI"] = ci; os.environ["GDPT_BVH_MAX_LEAF"] = ml; os.environ["GDPT_BVH_SPLIT"] = sp
    sc.close()
    g.gradient_path_render(sc, precision="f32", spp=2)
    best = None
    for _ in range(3):
        r = g.gradient_path_render(sc, precision="f32", spp=8)
        st = r["stats"]
        if best is None or st["ms_trace"] < best["ms_trace"]: best = st
    cnt = np.zeros(3)
    p = _params(sc, spp=1, precision="f32")
    _capi.check(_capi.lib().gdpt_traversal_counts(sc.context(0), C.byref(p), _capi.as_dp(cnt)))
    print("split %-4s ci %-4s max_leaf %s  trace %7.2f ms  shade %6.2f  nodes/ray %5.2f  prims/ray %5.2f  mean %.6f" % (sp, ci, ml, best["ms_trace"], best["ms_shade"], cnt[0], cnt[1], float(r.primal.mean())), flush=True)
