"""Parity gates for the BENCHMARKED mode and the BASELINE configs (round 2; VERDICT r1 "Next round" item 1):

 * f32 (the bench's arithmetic mode) per-sample records on every record scene incl. Sponza and Disney, with stated
   quantile bounds against the compiled reference's double records;
 * the north-star convergence gate (relMSE against the reference's 16384-spp render within 5 % of the CPU path's at
   equal spp) on Sponza, veach_mi (glossy) and Disney small films, both precisions;
 * the FULL C1 / C2 / C3 containers (cbox_gdpt 512x512, veach_mi_gradpath 768x512, disney_gradpath 683x512) rendered at
   reduced spp against the reference's pixel-seeded images through 8x8-block sums;
 * Box / Tent / Gaussian(2) pixel filters: the inputs of the reference's own src/tests/filter.cpp:15-66 re-hosted on the
   device's sample_primary, plus ray and record parity on filter-variant scenes.
Fixtures: tests/golden/make_golden_r2.py (run where the reference exists).  Nothing here reads /root/reference."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu

BLOCK = 8


@pytest.fixture(scope="module")
def scenes(g):
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = g.Scene.named(name)
        return cache[name]
    yield get
    for s in cache.values():
        s.close()


def block_sums(img):
    h, w, c = img.shape
    hp, wp = -(-h // BLOCK) * BLOCK, -(-w // BLOCK) * BLOCK
    pad = np.zeros((hp, wp, c)); pad[:h, :w] = img
    return pad.reshape(hp // BLOCK, BLOCK, wp // BLOCK, BLOCK, c).sum(axis=(1, 3))


# ------------------------------------------------------------------ f32 records, every record scene
# (bound on the fraction of records with any field off by more than 2e-3 relative, bound on the median error)
# measured on B200 (profiles/r2_f32_records.txt): 0.0003 / 0.0007 / 0.0000 / 0.0043 / 0.0014, medians <= 1e-9
F32_BOUNDS = {"cbox_gdpt_64": (0.005, 1e-6), "cbox_gdpt_d2": (0.005, 1e-6), "veach_mi_small": (0.005, 1e-6),
              "sponza_small": (0.02, 1e-6), "disney_small": (0.01, 1e-6)}


@pytest.mark.parametrize("name", sorted(F32_BOUNDS))
def test_fp32_records_close_to_reference_all_scenes(g, scenes, name):
    z = golden("records_%s.npz" % name)
    rec = g.grad_samples(scenes(name), z["xs"], z["ys"], z["streams"], precision="f32")
    want = z["rec"]
    finite = np.isfinite(want).all(axis=1) & np.isfinite(rec).all(axis=1)
    assert finite.mean() > 0.95
    err = np.abs(rec[finite] - want[finite]) / (1e-3 + np.abs(want[finite]))
    frac_off, med = (err > 2e-3).any(axis=1).mean(), np.median(err)
    bound_off, bound_med = F32_BOUNDS[name]
    assert frac_off < bound_off and med < bound_med, (name, frac_off, med, np.quantile(err, [0.9, 0.99]))
    # the quantity the images are made of: radiance and contrib/prob, summed over the records, agree closely
    for col in (slice(0, 3),):
        a, b = rec[finite][:, col].sum(axis=0), want[finite][:, col].sum(axis=0)
        assert np.allclose(a, b, rtol=2e-2, atol=1e-6), (name, a, b)


# ------------------------------------------------------------------ convergence gate on the other scene classes
def _relmse(i, r):
    ok = np.isfinite(i) & np.isfinite(r)
    return float(np.mean((i[ok] - r[ok]) ** 2 / (r[ok] ** 2 + 1e-3)))


@pytest.mark.parametrize("name", ["veach_mi_small", "sponza_small", "disney_small"])
def test_convergence_relmse_matches_the_cpu_path_other_scenes(g, scenes, name):
    """North-star gate 3 beyond the Cornell box: textured Sponza (the bench scene's geometry, materials, light), glossy
    veach_mi and the Disney BSDF.  Same seeds and streams as the CPU rows of the fixture; relMSE over the values finite
    in both images (the reference's Disney lobes emit NaN records)."""
    z = golden("relmse_%s.npz" % name)
    sc = scenes(name)
    spp = int(z["spp"])
    R = [z["R_primal"], z["R_dx"], z["R_dy"], z["R_poisson"]]
    cpu = z["cpu_relmse"].mean(axis=0)
    for prec in ("f64", "f32"):
        rows = []
        for sd in z["seeds"]:
            r = g.gradient_path_render(sc, spp=spp, precision=prec, seed=int(sd))
            rows.append([_relmse(i, ref) for i, ref in zip((r.primal, r.dx, r.dy, r.poisson), R)])
        gpu = np.mean(rows, axis=0)
        # a NaN record of the reference's Disney lobes turns its whole Poisson image into NaN (fourierSolve is global): that
        # column of the fixture is then NaN and there is nothing to compare
        ok = np.isfinite(cpu)
        assert ok[:3].all() and np.all(np.isfinite(gpu[ok]))
        assert np.all(np.abs(gpu[ok] / cpu[ok] - 1.0) < 0.05), (name, prec, gpu, cpu)


# ------------------------------------------------------------------ full-size C1 / C2 / C3
@pytest.mark.parametrize("name", ["cbox_gdpt", "veach_mi_gradpath", "disney_gradpath"])
def test_full_size_configs_against_reference_block_sums(g, scenes, name):
    """BASELINE configs C1-C3 at their full film size, reduced spp, pixel-seeded like the fixture: every 8x8 block of
    primal / dx / dy sums to what the compiled reference's images sum to (FP64 mode: equal but for the few samples
    where a last-ulp difference flips a branch; f32: close)."""
    z = golden("config_%s.npz" % name)
    sc = scenes(name)
    assert (sc.width, sc.height) == tuple(int(v) for v in z["size"])
    spp = int(z["spp"])
    # Disney: the reference's images hold NaN/Inf pixels (pow of negative bases; not the same ones on every run), and its
    # offsets read a stale std::optional at the second bounce ("Undefined behaviour") — in a scene with one or
    # two materials the stale vertex usually carries the base material, so a percent or two of its samples keep an offset
    # alive on garbage where the product drops it.  One such sample changes its 8x8 block: the gradient images are
    # therefore judged by "a good share of the blocks is identical and the rest correlates", the primal by the bound below.
    disney = name == "disney_gradpath"
    loose = 0.12 if disney else 0.01
    for prec, tol, frac in (("f64", 1e-7, loose), ("f32", 2e-2, max(loose, 0.05))):
        r = g.gradient_path_render(sc, spp=spp, precision=prec)
        for key, img in (("c", r.primal), ("cx", r.dx), ("cy", r.dy)):
            got = block_sums(np.nan_to_num(img, nan=0.0, posinf=0.0, neginf=0.0))
            want, scale = z["bs_" + key], z["abs_" + key]
            bad = np.abs(got - want) > tol * (scale + 1e-9)
            if disney and key != "c":
                corr = np.corrcoef(got.reshape(-1), want.reshape(-1))[0, 1]
                assert bad.mean() <= 0.85 and corr > 0.5, (name, prec, key, bad.mean(), corr)
                continue
            assert bad.mean() <= frac, (name, prec, key, bad.mean())
        if name != "disney_gradpath":
            # (a handful of samples may differ — depth ties decided in float, RR branch flips — and one bright sample moves the mean)
            assert np.allclose(r.primal.mean(axis=(0, 1)), z["mean_c"], rtol=5e-4 if prec == "f64" else 5e-3)
        assert np.isfinite(r.poisson).all() or name == "disney_gradpath"


# ------------------------------------------------------------------ pixel filters
FILTER_SCENES = {"cbox_gdpt_64_box": ("box", 2.0), "cbox_gdpt_64_tent": ("tent", 2.0), "cbox_gdpt_64_gauss2": ("gaussian", 2.0)}


@pytest.mark.parametrize("name", sorted(FILTER_SCENES))
def test_pixel_filters_match_reference(g, scenes, name):
    z = golden("filter_%s.npz" % name)
    sc = scenes(name)
    rays = g.sample_primary(sc, z["screen_pos"], precision="f64")
    assert np.abs(rays - z["rays"]).max() < 1e-12
    rays32 = g.sample_primary(sc, z["screen_pos"], precision="f32")
    # float Box-Muller: the error scales with the offset, up to ~10 pixels for stddev 2 against ~2.5 for the scenes' own 0.5
    assert np.abs(rays32[:, 3:] - z["rays"][:, 3:6]).max() < (8e-5 if name.endswith("gauss2") else 2e-5)
    rec = g.grad_samples(sc, z["xs"], z["ys"], z["streams"], precision="f64")
    want = z["rec"]
    finite = np.isfinite(want).all(axis=1)
    err = np.abs(rec[finite] - want[finite]) / (1e-12 + np.abs(want[finite]))
    diff = err > 1e-7
    # DESIGN.md 2, "Undefined behaviour": an offset still enabled at the second bounce reads a stale std::optional in the
    # reference; where the product dropped such an offset (contribS 0, w 1) and the reference found something else in its
    # stack, that offset's fields are not compared (they are counted) — the wide Gaussian enables more offsets than sigma 0.5
    ub = 0
    for c0, wi in ((6, 19), (12, 20), (9, 21), (15, 22)):
        got = rec[finite]
        dropped = (got[:, c0:c0 + 3] == 0).all(axis=1) & (got[:, wi] == 1) & (diff[:, c0:c0 + 3].any(axis=1) | diff[:, wi])
        diff[dropped, c0:c0 + 3] = False; diff[dropped, wi] = False
        ub += int(dropped.sum())
    assert diff.any(axis=1).mean() <= 4e-3 and np.median(err) < 1e-12 and ub <= 0.02 * finite.sum(), (diff.any(axis=1).mean(), ub)


def filter_offsets_from_rays(sc, dirs, px, py):
    """Invert camera.cpp:23-47 for rays through pixel (px, py): world direction -> camera direction -> the screen
    position whose sample_to_cam image lies on that direction -> offset = remapped_pos * (w, h) - (pixel + 0.5)."""
    w, h = sc.width, sc.height
    c2w = np.array(sc.a["camera.cam_to_world"], dtype=np.float64).reshape(4, 4)
    s2c = np.array(sc.a["camera.sample_to_cam"], dtype=np.float64).reshape(4, 4)
    dc = dirs @ np.linalg.inv(c2w[:3, :3]).T
    out = []
    for q in dc[:, :2] / dc[:, 2:3]:
        A = np.array([[s2c[0, 0] - q[0] * s2c[2, 0], s2c[0, 1] - q[0] * s2c[2, 1]], [s2c[1, 0] - q[1] * s2c[2, 0], s2c[1, 1] - q[1] * s2c[2, 1]]])
        b = np.array([q[0] * s2c[2, 3] - s2c[0, 3], q[1] * s2c[2, 3] - s2c[1, 3]])
        sx, sy = np.linalg.solve(A, b)
        out.append([sx * w - (px + 0.5), sy * h - (py + 0.5)])
    return np.array(out)


def filter_inverse_kernel(kind, width, s0):
    """1 / kernel at the sampled offset s0, as src/tests/filter.cpp:27-63 writes it"""
    if kind == "box":
        return width * width
    if kind == "tent":
        hh = width / 2
        return 1.0 / (((1 - abs(s0[0]) / hh) / hh) * ((1 - abs(s0[1]) / hh) / hh))
    return 1.0 / (np.exp(-((s0[0] ** 2 + s0[1] ** 2) / (width * width)) / 2) / (width * width * 2 * np.pi))


@pytest.mark.parametrize("name", sorted(FILTER_SCENES))
def test_filter_jacobian_known_answer_on_the_device(g, scenes, name):
    """src/tests/filter.cpp:15-66 re-hosted on the device: the finite-difference Jacobian of sample(filter, rnd) at
    rnd = (0.3, 0.4), eps = 1e-6, equals 1 / kernel within 1e-3 for Box(2), Tent(2), Gaussian(2).  The DEVICE's
    sample_primary does the sampling (rnd is the fractional part of the pixel position, camera.cpp:29-33); the filter
    offsets are read back out of its rays by inverting the pinhole map."""
    kind, width = FILTER_SCENES[name]
    sc = scenes(name)
    w, h = sc.width, sc.height
    eps, px, py = 1e-6, 20, 30
    sp = np.array([[(px + 0.3) / w, (py + 0.4) / h], [(px + 0.3 + eps) / w, (py + 0.4) / h], [(px + 0.3) / w, (py + 0.4 + eps) / h]])
    rays = g.sample_primary(sc, sp, precision="f64")
    off = filter_offsets_from_rays(sc, rays[:, 3:], px, py)
    ddu, ddv = (off[1] - off[0]) / eps, (off[2] - off[0]) / eps
    det = abs(ddu[0] * ddv[1] - ddu[1] * ddv[0])
    inv_kernel = filter_inverse_kernel(kind, width, off[0])
    assert abs(det - inv_kernel) < 1e-3 * max(1.0, inv_kernel), (det, inv_kernel)


# ------------------------------------------------------------------ wave scheduling
def test_waves_launched_without_host_syncs_give_the_same_images(g):
    """The default schedule (first wave read back by the host, the rest launched from its path counts with device-side
    counts, two waves in flight on two streams) against the fully host-synchronised loop: the same samples, so the same
    images up to the order of the floating-point accumulation."""
    imgs = {}
    for tag, opts in (("sync", {"async_waves": 0}), ("async1", {"async_waves": 1, "wave_streams": 1}), ("async2", {"async_waves": 1, "wave_streams": 2})):
        sc = g.Scene.named("cbox_gdpt")
        for k, v in opts.items():
            sc.set_option(k, v)
        imgs[tag] = [g.gradient_path_render(sc, spp=9, precision=p, wave_samples=2) for p in ("f64", "f32")]
        # a second render on the same context reuses the calibration and must not depend on what the streams did before
        again = g.gradient_path_render(sc, spp=9, precision="f32", wave_samples=2)
        assert np.allclose(again.primal, imgs[tag][1].primal, rtol=1e-9, atol=1e-12)
        st = imgs[tag][1].stats
        assert st["rays_traced"] > 9 * 512 * 512 * 5
        sc.close()
    for tag in ("async1", "async2"):
        for a, b in zip(imgs["sync"], imgs[tag]):
            for f in ("primal", "dx", "dy"):
                assert np.allclose(getattr(a, f), getattr(b, f), rtol=1e-9, atol=1e-12), (tag, f)
            assert np.linalg.norm(a.poisson - b.poisson) <= 1e-9 * np.linalg.norm(a.poisson)


def test_images_are_bit_reproducible(g):
    """Deterministic accumulation (per-path result records added in sample order, waves reduced in wave order): two renders
    give the same bits — also across wave schedules and wave sizes — and the atomicAdd(double) path agrees to rounding."""
    imgs = {}
    for tag, opts, wave in (("a", {}, 2), ("b", {}, 2), ("sync", {"async_waves": 0}, 2), ("one_stream", {"wave_streams": 1}, 2), ("waves_of_3", {}, 3),
                            ("atomics", {"deterministic": 0}, 2)):
        sc = g.Scene.named("cbox_gdpt")
        for k, v in opts.items():
            sc.set_option(k, v)
        imgs[tag] = g.gradient_path_render(sc, spp=9, precision="f32", wave_samples=wave)
        sc.close()
    for tag in ("b", "sync", "one_stream", "waves_of_3"):
        for f in ("primal", "dx", "dy", "poisson"):
            assert np.array_equal(getattr(imgs["a"], f), getattr(imgs[tag], f)), (tag, f)
    for f in ("primal", "dx", "dy"):
        assert np.allclose(getattr(imgs["a"], f), getattr(imgs["atomics"], f), rtol=1e-9, atol=1e-12), f


def test_path_and_shift_mode_images_are_bit_reproducible_and_equal_the_atomic_accumulation(g):
    """The `path` integrator and the shift-mapped gradpath mode accumulate through per-path result records and an ordered
    reduction (as the reference mode does): two renders give the same bits, several waves included, and the images equal the
    atomicAdd accumulation (`deterministic` = 0) up to its summation order.  Row bands compose (border rule of the shift mode
    at row 0 / column 0 included)."""
    sc = g.Scene.named("cbox_gdpt_64")
    for kw in (dict(spp=6, precision="f32", seed=5), dict(spp=5, precision="f64", seed=2, wave_samples=2)):
        a = g.path_render(sc, **kw); b = g.path_render(sc, **kw)
        assert np.array_equal(a, b)
        s1 = g.gradient_path_render(sc, offsets="shift", **kw); s2 = g.gradient_path_render(sc, offsets="shift", **kw)
        for n in ("primal", "dx", "dy", "poisson"):
            assert np.array_equal(s1[n], s2[n]), n
        sc.set_option("deterministic", 0)
        pa = g.path_render(sc, **kw); sa = g.gradient_path_render(sc, offsets="shift", **kw)
        sc.set_option("deterministic", 1)
        assert np.allclose(a, pa, rtol=1e-12, atol=1e-15)
        for n in ("primal", "dx", "dy"):
            assert np.allclose(s1[n], sa[n], rtol=1e-11, atol=1e-14), n
    sc.close()


def test_packed_triangle_records_give_the_same_bits_as_the_vertex_pools(g):
    """A hit triangle's vertex data comes from one packed record (96 bytes / three 256-bit loads in f32, 192 / six in f64) instead
    of the index and vertex pools — the same values, so the images are the same bits (textured mesh with normals and uvs, the
    Cornell box, sphere lights + rough plastic)."""
    for name in ("sponza_small", "cbox_gdpt_64", "veach_mi_small"):
        sc = g.Scene.named(name)
        a = g.gradient_path_render(sc, spp=4, precision="f32", seed=9)
        sc.set_option("tri_records", 0)
        b = g.gradient_path_render(sc, spp=4, precision="f32", seed=9)
        sc.set_option("tri_records", 1)
        p = g.path_render(sc, spp=2, precision="f32", seed=1) if name != "sponza_small" else None
        sc.set_option("tri_records", 0)
        q = g.path_render(sc, spp=2, precision="f32", seed=1) if name != "sponza_small" else None
        sc.close()
        for n in ("primal", "dx", "dy", "poisson"):
            assert np.array_equal(a[n], b[n]), (name, n)
        if p is not None:
            assert np.array_equal(p, q), name
    # the f64 mode has its own records (192 bytes, six loads)
    sc = g.Scene.named("sponza_small")
    a = g.gradient_path_render(sc, spp=2, precision="f64", seed=4)
    sc.set_option("tri_records", 0)
    b = g.gradient_path_render(sc, spp=2, precision="f64", seed=4)
    sc.close()
    for n in ("primal", "dx", "dy", "poisson"):
        assert np.array_equal(a[n], b[n]), ("f64", n)
