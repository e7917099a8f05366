"""CPU suite: host logic and the C-ABI surface (no compute calls without a GPU)."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT


def test_library_loads_and_exports_every_declared_symbol(g):
    hdr = open(os.path.join(ROOT, "include", "gdpt.h")).read()
    declared = set(re.findall(r"\b(gdpt_[a-z_0-9]+)\s*\(", hdr))
    declared -= {"gdpt_ctx"}
    assert declared == set(g._capi.EXPORTS), declared ^ set(g._capi.EXPORTS)
    lib = g._capi.lib()
    for name in declared:
        assert getattr(lib, name) is not None


def test_abi_struct_sizes_match_header(g, tmp_path):
    """ctypes mirrors must have the sizes a C compiler gives the header's structs."""
    import subprocess
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "gdpt.h"\nint main(void){printf("%zu %zu %zu %zu %zu\\n", sizeof(gdpt_render_params),'
                   ' sizeof(gdpt_images), sizeof(gdpt_stats), sizeof(gdpt_image), sizeof(gdpt_scene_desc));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    sizes = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    c = g._capi
    assert sizes == [ctypes.sizeof(t) for t in (c.GdptRenderParams, c.GdptImages, c.GdptStats, c.GdptImage, c.GdptSceneDesc)]


def test_no_cpu_fallback_without_device(g):
    lib = g._capi.lib()
    if lib.gdpt_device_count() > 0:
        pytest.skip("a CUDA device is present")
    u = np.zeros((4, 4, 3))
    with pytest.raises(g._capi.GdptError, match="no CUDA device"):
        g.fourier_solve(u, u, u)
    sc = g.Scene.named("cbox_gdpt_64")
    with pytest.raises(g._capi.GdptError, match="no CUDA device"):
        g.render(sc)
    # the entry points added in round 2 have no CPU path either; their argument checks answer before the device is looked for
    from gradient_based_path_tracing_b200 import smallgdpt
    with pytest.raises(g._capi.GdptError, match="no CUDA device"):
        g.reconstruct_l1(u, u, u)
    with pytest.raises(g._capi.GdptError, match="no CUDA device"):
        smallgdpt.render(64, 48, 1)
    with pytest.raises(g._capi.GdptError, match="outside 2..16384"):
        smallgdpt.render(1, 48, 1)
    with pytest.raises(g._capi.GdptError, match="2x2"):
        g.reconstruct_l1(u[:1], u[:1], u[:1])


def test_struct_layout_of_the_l1_parameters_matches_the_header(g, tmp_path):
    import ctypes
    import subprocess
    src = tmp_path / "l1.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "gdpt.h"\nint main(void){printf("%zu %zu %zu %zu\\n", sizeof(gdpt_l1_params),'
                   ' offsetof(gdpt_l1_params, reweight_primal), offsetof(gdpt_l1_params, alpha), offsetof(gdpt_l1_params, cg_tolerance));return 0;}\n')
    exe = tmp_path / "l1"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    P = g._capi.GdptL1Params
    assert got == [ctypes.sizeof(P), P.reweight_primal.offset, P.alpha.offset, P.cg_tolerance.offset]


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "gradient_based_path_tracing_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "liboracle" not in txt and "lajolla_ref" not in txt, f


def test_container_roundtrip(tmp_path, g):
    from gradient_based_path_tracing_b200.scene import read_container, write_container
    arrays = {"a": np.arange(12, dtype=np.float64).reshape(3, 4), "b.c": np.array([1, 2, 3], dtype=np.int32),
              "e": np.zeros((0, 3), dtype=np.float64), "u": np.arange(5, dtype=np.uint8)}
    for name in ("x.gdpt", "x.gdpt.z"):
        p = str(tmp_path / name)
        write_container(p, arrays)
        back = read_container(p)
        assert set(back) == set(arrays)
        for k in arrays:
            assert back[k].dtype == arrays[k].dtype and np.array_equal(back[k], arrays[k])


@pytest.mark.parametrize("name,size,spp", [("cbox_gdpt", (512, 512), 4), ("veach_mi_gradpath", (768, 512), 64),
                                           ("disney_gradpath", (683, 512), 256), ("sponza_1080p", (1920, 1080), 512),
                                           ("sponza_2160p", (3840, 2160), 1024)])
def test_shipped_scene_containers_match_baseline_configs(g, name, size, spp):
    sc = g.Scene.named(name)
    assert (sc.width, sc.height) == size
    assert sc.options.samples_per_pixel == spp
    assert sc.options.integrator == g.Integrator.GradPath and int(g.Integrator.GradPath) == 7   # src/scene.h:14-23
    d = sc.desc()
    assert d.n_shapes == sc.a["shapes"].shape[0] and d.n_materials == sc.a["materials"].shape[0]
    if name.startswith("sponza"):
        assert d.n_triangles == 66445 and d.n_shapes == 38 and d.n_images3 == 10    # 37 meshes + 1 sphere light


def test_render_rejects_other_integrators(g):
    sc = g.Scene.named("cbox_gdpt_64")
    sc.options.integrator = g.Integrator.VolPath       # Path and GradPath are built; the rest ends like render.cpp:390-393
    with pytest.raises(AssertionError):
        g.render(sc)


def test_row_band_partition(g):
    from gradient_based_path_tracing_b200.multi import partition_rows
    for h in (1, 3, 64, 90, 575, 1080, 2160):
        for n in (1, 2, 3, 4, 8):
            bands = partition_rows(h, n)
            assert len(bands) == n
            assert bands[0][0] == 0 and bands[-1][1] == h
            for (a0, a1), (b0, b1) in zip(bands, bands[1:]):
                assert a1 == b0 and a0 <= a1
            assert all(b0 % 4 == 0 for b0, b1 in bands if b1 > b0)   # non-empty bands start on a 4-row tile boundary
            sizes = [b1 - b0 for b0, b1 in bands if b1 > b0]
            if h >= 4 * n:
                assert max(sizes) - min(sizes) <= 4 + h % 4


# ------------------------------------------------------------------ host BVH builder (product code, CPU-only part)
def _bvh_check(tmp_path, scene_name, env=None):
    import json
    import subprocess
    import gradient_based_path_tracing_b200 as g
    exe = tmp_path / "bvh_check"
    if not exe.exists():
        csrc = os.path.join(ROOT, "gradient_based_path_tracing_b200", "csrc")
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", csrc, os.path.join(ROOT, "tests", "host", "bvh_check.cpp"),
                               os.path.join(csrc, "gdpt_bvh.cpp"), "-o", str(exe), "-lpthread"])
    a = g.Scene.named(scene_name).a
    files = []
    for key, dt in (("shapes", np.int32), ("shape_offsets", np.int64), ("spheres", np.float64), ("positions", np.float64), ("indices", np.int32)):
        f = tmp_path / (scene_name + "." + key + ".bin")
        np.ascontiguousarray(a[key], dtype=dt).tofile(f)
        files.append(str(f))
    e = dict(os.environ)
    e.update(env or {})
    return json.loads(subprocess.check_output([str(exe)] + files, env=e, timeout=300)), a


@pytest.mark.parametrize("scene_name", ["cbox_gdpt", "veach_mi_gradpath", "sponza_1080p"])
def test_bvh_builder_invariants(tmp_path, scene_name):
    """Every primitive reachable, leaf ranges valid, depth within the traversal stack, bounds = float min/max of the geometry
    (what rtcGetSceneBounds reports), two builds identical (the parallel subtree build must stay deterministic)."""
    r, a = _bvh_check(tmp_path, scene_name)
    assert r["bad_range"] == 0 and r["bad_child"] == 0 and r["unvisited"] == 0 and r["uncovered"] == 0
    assert r["distinct_prims"] == r["scene_prims"]
    assert r["prim_records"] >= r["scene_prims"] and r["max_leaf"] <= 7
    assert r["depth"] == r["reported_depth"] and 3 * r["depth"] + 1 <= 64
    assert r["deterministic"] is True
    tri = a["shapes"][:, 0] == 1
    if tri.all():
        p = np.asarray(a["positions"], dtype=np.float64).astype(np.float32)
        assert np.array_equal(np.float32(r["lo"]), p.min(axis=0)) and np.array_equal(np.float32(r["hi"]), p.max(axis=0))


def test_bvh_presplit_only_adds_references(tmp_path):
    r0, _ = _bvh_check(tmp_path, "sponza_1080p", {"GDPT_BVH_SPLIT": "0"})
    r2, _ = _bvh_check(tmp_path, "sponza_1080p", {"GDPT_BVH_SPLIT": "2"})
    assert r0["prim_records"] == r0["scene_prims"] == r2["scene_prims"] == r2["distinct_prims"]
    assert r0["scene_prims"] < r2["prim_records"] <= 3 * r0["scene_prims"] + 1
    assert r0["lo"] == r2["lo"] and r0["hi"] == r2["hi"]


# ------------------------------------------------------------------ C++ host (lajolla_b200): container reader + flattening
def _fnv(buf, h=1469598103934665603):
    # same 64-bit FNV-1a as lajolla_b200 -describe (vectorised per byte would be slow in Python: hash via numpy loop-free trick is
    # not possible for FNV, so hash a strided sample of large arrays the same way on both sides is not an option either) -> plain loop on bytes
    for b in bytes(buf):
        h = ((h ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


@pytest.mark.parametrize("name,overlay", [("cbox_gdpt_64", None), ("veach_mi_small", None), ("disney_small", "disney_gradpath")])
def test_cpp_host_flattens_scenes_like_the_python_host(name, overlay):
    """lajolla_b200 (C++ host over the C ABI) must hand gdpt_create the same scene as scene.py does."""
    import json
    import subprocess
    import gradient_based_path_tracing_b200 as g
    exe = os.path.join(ROOT, "gradient_based_path_tracing_b200", "lajolla_b200")
    if not os.path.exists(exe):
        pytest.skip("lajolla_b200 not built")
    sdir = os.path.join(ROOT, "scenes")
    if overlay:
        cmd = [exe, "-describe", "-overlay", os.path.join(sdir, name + ".camera.gdpt"), os.path.join(sdir, overlay + ".gdpt.z")]
    else:
        cmd = [exe, "-describe", os.path.join(sdir, name + ".gdpt.z")]
    got = json.loads(subprocess.check_output(cmd, timeout=120).decode().strip().splitlines()[-1])
    sc = g.Scene.named(name)
    d = sc.desc()
    for k, v in (("width", d.width), ("height", d.height), ("filter_kind", d.filter_kind), ("filter_param", d.filter_param),
                 ("spp", d.samples_per_pixel), ("max_depth", d.max_depth), ("rr_depth", d.rr_depth), ("integrator", int(sc.options.integrator)),
                 ("n_shapes", d.n_shapes), ("n_positions", d.n_positions), ("n_triangles", d.n_triangles), ("n_normals", d.n_normals),
                 ("n_uvs", d.n_uvs), ("n_materials", d.n_materials), ("n_textures", d.n_textures), ("n_images3", d.n_images3),
                 ("n_images1", d.n_images1), ("n_lights", d.n_lights)):
        assert got[k] == v, k
    a = sc.a
    f8 = lambda k: np.ascontiguousarray(a[k], dtype=np.float64)
    assert got["h_camera"] == _fnv(f8("camera.cam_to_world"), _fnv(f8("camera.sample_to_cam")))
    assert got["h_shapes"] == _fnv(np.ascontiguousarray(a["shapes"], dtype=np.int32))
    assert got["h_offsets"] == _fnv(np.ascontiguousarray(a["shape_offsets"], dtype=np.int64))
    assert got["h_spheres"] == _fnv(f8("spheres"))
    assert got["h_positions"] == _fnv(f8("positions"))
    assert got["h_indices"] == _fnv(np.ascontiguousarray(a["indices"], dtype=np.int32))
    assert got["h_materials"] == _fnv(f8("materials")) and got["h_textures"] == _fnv(f8("textures")) and got["h_lights"] == _fnv(f8("lights"))
    himg = 1469598103934665603
    for prefix in ("img3", "img1"):
        for i in range(sc.n_images(prefix)):
            base = "%s.%d" % (prefix, i)
            px = f8(base + ".data") if base + ".data" in a else f8(base + ".palette")[np.asarray(a[base + ".index"])]
            himg = _fnv(np.ascontiguousarray(px, dtype=np.float64), himg)
    assert got["h_images"] == himg


def test_cpp_host_rejects_bad_input(tmp_path):
    import subprocess
    exe = os.path.join(ROOT, "gradient_based_path_tracing_b200", "lajolla_b200")
    if not os.path.exists(exe):
        pytest.skip("lajolla_b200 not built")
    bad = tmp_path / "bad.gdpt"
    bad.write_bytes(b"not a container")
    r = subprocess.run([exe, "-describe", str(bad)], capture_output=True, timeout=60)
    assert r.returncode == 1 and b"not a GDPT1 container" in r.stderr
    r = subprocess.run([exe, "-describe", str(tmp_path / "missing.gdpt")], capture_output=True, timeout=60)
    assert r.returncode == 1 and b"cannot open" in r.stderr


# ------------------------------------------------------------------ image writers of the C++ host (imwrite, image.cpp:134-173)
def _read_exr(path):
    """Independent minimal OpenEXR reader (scan lines, NONE / ZIP, HALF channels) for the writer's round trip."""
    import struct
    import zlib
    b = open(path, "rb").read()
    assert b[:4] == bytes([0x76, 0x2f, 0x31, 0x01]) and b[4] == 2
    p, attrs = 8, {}
    while b[p] != 0:
        e = b.index(0, p); name = b[p:e].decode(); p = e + 1
        e = b.index(0, p); typ = b[p:e].decode(); p = e + 1
        size = struct.unpack_from("<i", b, p)[0]; p += 4
        attrs[name] = (typ, b[p:p + size]); p += size
    p += 1
    chans, q, cl = [], 0, attrs["channels"][1]
    while cl[q] != 0:
        e = cl.index(0, q); chans.append((cl[q:e].decode(), struct.unpack_from("<i", cl, e + 1)[0])); q = e + 1 + 16
    assert [c[0] for c in chans] == ["B", "G", "R"] and all(c[1] == 1 for c in chans)
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    comp = attrs["compression"][1][0]
    lines = {0: 1, 3: 16}[comp]
    nblocks = (h + lines - 1) // lines
    offs = struct.unpack_from("<%dQ" % nblocks, b, p)
    img = np.zeros((h, w, 3), dtype=np.float32)
    for o in offs:
        y, size = struct.unpack_from("<2i", b, o)
        data = b[o + 8:o + 8 + size]
        n = min(lines, h - y) * w * 3 * 2
        if comp == 3 and size < n:
            t = np.frombuffer(zlib.decompress(data), dtype=np.uint8).astype(np.int32)
            t = (np.cumsum(t - np.concatenate([[0], np.full(n - 1, 128)])) % 256).astype(np.uint8)   # undo the predictor
            raw = np.empty(n, dtype=np.uint8)
            raw[0::2], raw[1::2] = t[:(n + 1) // 2], t[(n + 1) // 2:]
        else:
            raw = np.frombuffer(data, dtype=np.uint8)
        px = raw.view(np.float16).reshape(-1, 3, w).astype(np.float32)      # [line][B,G,R][x]
        img[y:y + px.shape[0]] = px[:, ::-1, :].transpose(0, 2, 1)
    return img


@pytest.mark.parametrize("w,h", [(8, 5), (37, 41), (160, 90)])
def test_cpp_host_exr_and_pfm_writers(tmp_path, w, h):
    """write_image (csrc/host): the EXR the reference would write by default (fp16 B/G/R, ZIP; NONE below 16x16) read back by an
    independent reader, by OpenCV and by the reference's own imread3 (tinyexr) where the oracle is built; PFM as imwrite."""
    import subprocess
    csrc = os.path.join(ROOT, "gradient_based_path_tracing_b200", "csrc")
    exe = tmp_path / "exr_check"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", os.path.join(csrc, "host"), "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "host", "exr_check.cpp"), os.path.join(csrc, "host", "gdpt_container.cpp"), "-o", str(exe), "-lz"])
    rng = np.random.default_rng(w * h)
    img = rng.random((h, w, 3)) * np.array([1.0, 40.0, 1e-3]) - 0.1
    img[0, 0] = [0.0, 65504.0, 1e-8]            # zero, the largest half, a half denormal
    img[1, 1] = [1e6, -1e6, 6.1e-5]             # overflow to +-inf, smallest normal
    raw = tmp_path / "img.f64"
    img.tofile(raw)
    with np.errstate(over="ignore"):
        want = img.astype(np.float32).astype(np.float16).astype(np.float32)
    subprocess.check_call([str(exe), str(w), str(h), str(raw), str(tmp_path / "o.exr")])
    got = _read_exr(str(tmp_path / "o.exr"))
    assert np.array_equal(got, want)
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    try:
        import cv2
        cvimg = cv2.imread(str(tmp_path / "o.exr"), cv2.IMREAD_UNCHANGED)
    except Exception:
        cvimg = None
    if cvimg is not None:
        fin = np.isfinite(want)
        assert np.array_equal(cvimg[..., ::-1][fin], want[fin])
    from oracle import ref
    if ref.available():
        back = ref.imread3(str(tmp_path / "o.exr"))
        fin = np.isfinite(want)
        assert np.array_equal(back.astype(np.float32)[fin], want[fin])
    subprocess.check_call([str(exe), str(w), str(h), str(raw), str(tmp_path / "o.pfm")])
    b = open(tmp_path / "o.pfm", "rb").read()
    head = ("PF\n%d %d\n-1\n" % (w, h)).encode()
    assert b.startswith(head)
    assert np.array_equal(np.frombuffer(b[len(head):], dtype=np.float32).reshape(h, w, 3), img.astype(np.float32))


# ------------------------------------------------------------------ FFT form of the DCT-I (device functions compiled for the host)
def test_fft_dct1_device_functions_on_the_host(tmp_path):
    """csrc/gdpt_poisson_fft.cuh — the radix-4 forward / inverse FFT passes and the Bluestein DCT-I line transform the Poisson
    kernels run in shared memory — compiled for the host with one emulated thread (tests/host/fft_check.cpp) against the
    O(N^2) definition of FFTW's REDFT00, at the film sizes of the configs (incl. odd log2 P and the 4097 limit)."""
    import json
    import subprocess
    exe = tmp_path / "fft_check"
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "gradient_based_path_tracing_b200", "csrc"),
                           os.path.join(ROOT, "tests", "host", "fft_check.cpp"), "-o", str(exe)])
    sizes = [16, 17, 33, 64, 90, 512, 575, 768, 1080, 1920, 2160, 3840, 4097]
    r = json.loads(subprocess.check_output([str(exe)] + [str(n) for n in sizes], timeout=300))
    for n in sizes:
        assert r[str(n)]["roundtrip"] < 1e-14 and r[str(n)]["dct_rel_err"] < 1e-13, (n, r[str(n)])


def test_fft_line_buffer_swizzle_is_conflict_free_in_the_bank_model():
    """fft_sw (csrc/gdpt_poisson_fft.cuh) in the 8-bank / quarter-warp model of 16-byte shared-memory accesses: every access
    pattern of both FFT directions at P = 2048 / 4096 / 8192 costs the ideal number of wavefronts, where the natural layout
    costs 1.5-1.6 x (ncu measured 1.52 x on B200: profiles/r2_k_dct_cols_swizzled.txt).  The Python map is the header's."""
    import re
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import fft_swizzle_search as F
    src = open(os.path.join(ROOT, "gradient_based_path_tracing_b200", "csrc", "gdpt_poisson_fft.cuh")).read()
    assert re.search(r"fft_sw\(int e\) \{ const int x = \(e >> 3\) & 3; return e \^ \(x \^ \(x << 1\)\); \}", src)
    for P, logP in F.SIZES:
        natural, _ = F.cost(lambda e: e, P, logP)
        swizzled, worst = F.cost(F.fft_sw, P, logP)
        assert swizzled == 1.0, (P, worst)
        assert natural > 1.45
        assert sorted(F.fft_sw(e) for e in range(P)) == list(range(P))          # a permutation of the line


# ------------------------------------------------------------------ bench.py: the reference arm and the helpers that need no GPU
def test_bench_reference_arm_line_and_affine_cost_model():
    """`bench.py --impl reference` on a small film: one JSON line with the contract's keys, the affine cost model
    (fixed_s + spp * per_spp_s reproduces ms_per_step), the share of the stand-in ray caster, and the workload string the
    native arm prints for the same configuration."""
    import json
    import subprocess
    import sys
    from oracle import ref
    if not ref.scenes_available():
        pytest.skip("oracle/_ref not built")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--scene", "cbox_gdpt_64", "--steps", "2", "--warmup", "1",
                          "--ref-total-seconds", "3", "--cpu-seconds", "1"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config", "impl", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["dtype"] == "f64" and d["vs_baseline"] is None and d["e2e"]["value"] == d["value"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and 0.0 <= cb["rtc_fraction"] <= 1.0
    assert abs((cb["fixed_s"] + 4 * cb["per_spp_s"]) * 1e3 - d["ms_per_step"]) < 1e-6 * d["ms_per_step"] + 1e-9        # cbox_gdpt_64: 4 spp
    assert cb["value_if_ray_casting_were_free"] >= cb["value"]
    sys.path.insert(0, ROOT)
    import bench
    assert d["config"]["workload"] == bench.workload_name("cbox_gdpt_64", 64, 64, 4)
    m = bench.synthetic_camera(1)
    assert np.allclose(m[:3, :3].T @ m[:3, :3], np.eye(3), atol=1e-12) and np.allclose(m[3], [0, 0, 0, 1])
