"""small_gdpt.cpp (the reference's second program) — CPU side of its parity chain:

  reference program, compiled unmodified (oracle/_ref/smallgdpt_ref)  ->  tests/golden/smallgdpt_*.npz
  csrc/gdpt_smallgdpt.cuh compiled for the HOST (tests/host/smallgdpt_check.cpp)  ==  those fixtures, bit for bit

The GPU half (tests/test_smallgdpt_gpu.py) compares the same header running on the device with the same fixtures.
"""
import ctypes
import glob
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
CSRC = os.path.join(ROOT, "gradient_based_path_tracing_b200", "csrc")
GOLDEN = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "smallgdpt_*.npz")))
NAMES = ("image", "image_dx", "image_dy", "image_poisson")


@pytest.fixture(scope="module")
def check_exe(tmp_path_factory):
    exe = tmp_path_factory.mktemp("sg") / "smallgdpt_check"
    subprocess.check_call(["g++", "-O3", "-std=c++17", "-I", CSRC, os.path.join(ROOT, "tests", "host", "smallgdpt_check.cpp"), "-o", str(exe)])
    return str(exe)


def merged(five):
    """main()'s merge of the four one-sided gradient images (small_gdpt.cpp:530-539)"""
    c, cx0, cy0, cx1, cy1 = five
    cx = cx0.copy(); cx[:, 1:] += cx1[:, :-1]
    cy = cy0.copy(); cy[1:] += cy1[:-1]
    return c, cx, cy


def host_images(exe, w, h, spp, tmp_path):
    from oracle import cport
    out = tmp_path / ("sg_%dx%d_%d.bin" % (w, h, spp))
    subprocess.check_call([exe, str(w), str(h), str(spp), str(out)], timeout=600)
    c, cx, cy = merged(np.fromfile(out).reshape(5, h, w, 3))
    return {"image": c.astype(np.float32), "image_dx": np.abs(cx).astype(np.float32), "image_dy": np.abs(cy).astype(np.float32),
            "image_poisson": cport.poisson(c, cx, cy, 0.04).astype(np.float32)}


def test_fixtures_exist():
    assert len(GOLDEN) == 3


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_fixtures_are_what_the_compiled_reference_program_writes(path):
    from oracle import smallgdpt as ref
    if not ref.available():
        pytest.skip("oracle/_ref/smallgdpt_ref not built (needs /root/reference)")
    g = np.load(path)
    w, h = int(g["width"]), int(g["height"])
    keep = ~ref.stale_read_mask(w, h)
    for threads in (1, 3):
        # the rows are independent of the thread count, but the program reads stack words it never wrote (see
        # oracle/smallgdpt stale_read_mask): gradients are compared outside the pixels that can depend on them, the
        # reconstruction (which mixes all pixels) only when the gradients came out as in the fixture
        r = ref.run(w, h, int(g["spp"]), threads=threads)
        assert np.array_equal(r["image"], g["image"])
        for n in ("image_dx", "image_dy"):
            assert np.array_equal(r[n][keep], g[n][keep]), (n, threads)
        if all(np.array_equal(r[n], g[n]) for n in ("image_dx", "image_dy")):
            assert np.array_equal(r["image_poisson"], g["image_poisson"])


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p) for p in GOLDEN])
def test_device_functions_compiled_for_the_host_equal_the_reference_bit_for_bit(path, check_exe, tmp_path):
    g = np.load(path)
    mine = host_images(check_exe, int(g["width"]), int(g["height"]), int(g["spp"]), tmp_path)
    for n in ("image", "image_dx", "image_dy"):
        assert np.array_equal(mine[n], g[n]), (n, float(np.abs(mine[n] - g[n]).max()))
    # the solve here is the oracle's DCT (the program's own solve ran on the same stand-in): equal to float32 rounding
    np.testing.assert_allclose(mine["image_poisson"], g["image_poisson"], rtol=2e-6, atol=1e-7)


def test_erand48_stream_is_libc_s(check_exe):
    """sg::erand against glibc's erand48 for the row seeds of small_gdpt.cpp:462 (rng = {0, 0, y*y*y})"""
    libc = ctypes.CDLL(None)
    libc.erand48.restype = ctypes.c_double
    libc.erand48.argtypes = [ctypes.POINTER(ctypes.c_ushort)]
    for y in (0, 1, 7, 41, 300, 767, 2000):
        mine = [float.fromhex(v) for v in subprocess.check_output([check_exe, "erand", str(y), "64"], text=True).split()]
        st = (ctypes.c_ushort * 3)(0, 0, (y * y * y) & 0xFFFF)
        theirs = [libc.erand48(st) for _ in range(64)]
        assert mine == theirs, y


def test_the_header_s_two_stated_choices_matter(tmp_path):
    """gdpt_smallgdpt.cuh keeps two compiler-dependent behaviours of the reference program; the other reading of the
    argument order (origin from the un-normalised direction) is NOT what the compiled program does."""
    exe = tmp_path / "sgcheck0"
    subprocess.check_call(["g++", "-O3", "-std=c++17", "-DSG_ORIGIN_AFTER_NORM=0", "-I", CSRC, os.path.join(ROOT, "tests", "host", "smallgdpt_check.cpp"), "-o", str(exe)])
    g = np.load(GOLDEN[-1])
    w, h, spp = int(g["width"]), int(g["height"]), int(g["spp"])
    out = tmp_path / "o.bin"
    subprocess.check_call([str(exe), str(w), str(h), str(spp), str(out)], timeout=600)
    c = np.fromfile(out).reshape(5, h, w, 3)[0].astype(np.float32)
    assert not np.array_equal(c, g["image"])
    assert np.abs(c - g["image"]).max() < 1e-2            # a last-digits difference, not another image


def test_cli_host_fails_loudly_without_a_device(tmp_path):
    """smallgdpt_b200 has no CPU path: without a CUDA device it says so and exits non-zero (no files written)."""
    exe = os.path.join(ROOT, "gradient_based_path_tracing_b200", "smallgdpt_b200")
    if not os.path.exists(exe):
        pytest.skip("smallgdpt_b200 not built")
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a CUDA device is present")
    except ImportError:
        pass
    r = subprocess.run([exe, "1", "-size", "16", "12"], cwd=tmp_path, capture_output=True, timeout=60)
    assert r.returncode != 0 and b"no CUDA device" in r.stderr
    assert not list(tmp_path.glob("*.pfm"))
