"""Two 1-spp renders of the 1080p Sponza film: the workload under which the Poisson kernels are captured by ncu."""
import sys
sys.path.insert(0, '.')
import gradient_based_path_tracing_b200 as g
sc = g.Scene.named("sponza_1080p")
for _ in range(2):
    r = g.gradient_path_render(sc, spp=1, precision="f32", seed=3)
print(r.stats["ms_poisson"])
