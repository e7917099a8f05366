import sys, time, numpy as np
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
sc = g.Scene.named("sponza_1080p")
t=time.time(); sc.context(0); print("create", time.time()-t)
for prec in ("f32", "f64"):
    for spp in (2, 8):
        t=time.time(); r = g.gradient_path_render(sc, precision=prec, spp=spp); dt=time.time()-t
        st = r["stats"]
        print(prec, spp, "wall %.3f" % dt, "primal mean %.5f poisson mean %.5f" % (r.primal.mean(), r.poisson.mean()),
              "Msamples/s(acc) %.1f" % (1920*1080*spp/ (st['ms_total']-st['ms_poisson']-st['ms_copy'])/1e3), st)
