"""GPU probe: time of merge + screened-Poisson solve (stats.ms_poisson: CUDA events around k_merge .. the last solve kernel inside
gdpt_render) at the film sizes of the configs, FFT form against the dense-matrix form, and the difference of their results."""
import sys
sys.path.insert(0, '.')
import numpy as np
import gradient_based_path_tracing_b200 as g

for name in ["cbox_gdpt", "veach_mi_gradpath", "sponza_1080p", "sponza_2160p"]:
    res = {}
    for fft in (1, 0):
        sc = g.Scene.named(name)
        sc.set_option("poisson_fft", fft)
        g.gradient_path_render(sc, spp=1, precision="f32", seed=3)
        ms = []
        for _ in range(5):
            r = g.gradient_path_render(sc, spp=1, precision="f32", seed=3)
            ms.append(r.stats["ms_poisson"])
        res[fft] = (min(ms), r.poisson.copy())
        sc.close()
    rel = np.linalg.norm(res[1][1] - res[0][1]) / np.linalg.norm(res[0][1])
    h, w, _ = res[1][1].shape
    print("%-18s %4dx%-4d  FFT form %7.3f ms   matrix form %7.3f ms   rel-L2 between them %.2e   (96 B/px at %.1f GB/s)" % (
        name, w, h, res[1][0], res[0][0], rel, 96.0 * w * h / (res[1][0] * 1e-3) / 1e9), flush=True)
