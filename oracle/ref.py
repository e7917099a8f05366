"""TEST INFRASTRUCTURE — ctypes driver for the REAL reference built by oracle/build_ref.sh.

Only tests/, tools that generate golden fixtures, __graft_entry__.smoke() and
bench.py's CPU arm may import this.  The product never does.

Every call below lands in the reference's own code (see oracle/ref_shims/ref_hooks.cpp
for the file:line of each function reached).
"""
import ctypes as C
import os
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "liblajolla_ref.so")
REFERENCE_DIR = os.environ.get("GDPT_REFERENCE_DIR", "/root/reference")
DEFAULT_SEED = 0x31E241F862A1FB5E  # src/pcg.h:33

_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/liblajolla_ref.so missing: run oracle/build_ref.sh where /root/reference exists")
        L = C.CDLL(LIB_PATH)
        dp = C.POINTER(C.c_double)
        ip = C.POINTER(C.c_int)
        L.gdpt_ref_init.argtypes = [C.c_int]
        L.gdpt_ref_load.restype = C.c_void_p
        L.gdpt_ref_load.argtypes = [C.c_char_p]
        L.gdpt_ref_free.argtypes = [C.c_void_p]
        L.gdpt_ref_info.argtypes = [C.c_void_p, ip]
        L.gdpt_ref_set_options.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.gdpt_ref_render.argtypes = [C.c_void_p, dp, dp, dp, dp]
        L.gdpt_ref_render_path.argtypes = [C.c_void_p, dp]
        L.gdpt_ref_fourier_solve.argtypes = [C.c_int, C.c_int, dp, dp, dp, C.c_double, dp]
        L.gdpt_ref_grad_sample.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_uint64, dp]
        L.gdpt_ref_render_pixelseed.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_int, dp, dp, dp, dp, dp]
        L.gdpt_ref_sample_primary.argtypes = [C.c_void_p, C.c_int, dp, dp]
        L.gdpt_ref_intersect.argtypes = [C.c_void_p, C.c_int, dp, dp, ip, dp]
        L.gdpt_ref_bsdf.argtypes = [C.c_void_p, C.c_int, C.c_int, dp, dp, dp, dp]
        L.gdpt_ref_export.argtypes = [C.c_void_p, C.c_char_p]
        L.gdpt_ref_render_aux.argtypes = [C.c_void_p, C.c_int, dp]
        L.gdpt_ref_imread3.argtypes = [C.c_char_p, ip, ip, dp]
        L.gdpt_ref_path_sample.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_uint64, dp]
        L.gdpt_ref_shift_sample.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_uint64, dp]
        L.gdpt_ref_sample_light.argtypes = [C.c_void_p, C.c_int, dp, dp, dp]
        L.gdpt_ref_occluded.argtypes = [C.c_void_p, C.c_int, dp, ip]
        L.gdpt_ref_render_path_pixelseed.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_int, dp]
        L.gdpt_rtc_standin_stats.argtypes = [C.POINTER(C.c_uint64), C.c_int]
        L.gdpt_rtc_standin_tsc.restype = C.c_uint64
        _lib = L
    return _lib


class _quiet:
    """The reference prints a progress line per tile and banners to stdout (progress_reporter.h,
    parse_scene.cpp); silence fd 1 around its calls so callers keep a clean stdout (bench.py's JSON line)."""

    def __enter__(self):
        import sys
        sys.stdout.flush()
        self.saved = os.dup(1)
        self.null = os.open(os.devnull, os.O_WRONLY)
        os.dup2(self.null, 1)

    def __exit__(self, *a):
        try:
            C.CDLL(None).fflush(None)
        except Exception:
            pass
        os.dup2(self.saved, 1)
        os.close(self.saved)
        os.close(self.null)


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double)) if a is not None else None


_threads = None


def init(threads=None):
    global _threads
    threads = threads or os.cpu_count() or 1
    if _threads != threads:
        if _threads is None:
            import atexit
            atexit.register(lib().gdpt_ref_shutdown)  # worker threads must be joined before static destructors run
        lib().gdpt_ref_init(threads)
        _threads = threads
    return threads


def rtc_stats(reset=False):
    """(TSC ticks inside the stand-in's rtcIntersect1/rtcOccluded1 summed over threads, calls)."""
    out = (C.c_uint64 * 2)()
    lib().gdpt_rtc_standin_stats(out, 1 if reset else 0)
    return int(out[0]), int(out[1])


def tsc():
    return int(lib().gdpt_rtc_standin_tsc())


def make_variant(xml_path, out_dir=None, **subs):
    """Write a variant of a reference scene XML next to symlinks of its assets.

    subs: integrator='gradpath', width=, height=, spp=, max_depth=, extra_xml= (text
    spliced before </scene>).  The reference tree is never modified.
    """
    import re
    src_dir = os.path.dirname(os.path.abspath(xml_path))
    out_dir = out_dir or tempfile.mkdtemp(prefix="gdpt_scene_")
    os.makedirs(out_dir, exist_ok=True)
    for e in os.listdir(src_dir):
        d = os.path.join(out_dir, e)
        if not os.path.lexists(d) and not e.endswith(".xml"):
            os.symlink(os.path.join(src_dir, e), d)
    txt = open(xml_path).read()
    if "integrator" in subs:
        txt = re.sub(r'<integrator\s+type="[^"]*"', '<integrator type="%s"' % subs["integrator"], txt, count=1)
    if "max_depth" in subs:
        if re.search(r'name="maxDepth"', txt):
            txt = re.sub(r'(name="maxDepth"\s+value=")[^"]*"', r'\g<1>%d"' % subs["max_depth"], txt)
        else:
            txt = re.sub(r'(<integrator[^>]*)/>', r'\1>\n<integer name="maxDepth" value="%d"/></integrator>' % subs["max_depth"], txt, count=1)
    for key, name in (("width", "width"), ("height", "height"), ("spp", "sampleCount")):
        if key in subs:
            txt = re.sub(r'(name="%s"\s+value=")[^"]*"' % name, r'\g<1>%d"' % subs[key], txt)
    if "rfilter" in subs:      # (type, parameter name, value): replaces the film's reconstruction filter (parse_scene.cpp:620-650)
        t, n, v = subs["rfilter"]
        txt, k = re.subn(r'<rfilter\s+type="[^"]*"\s*/>', '<rfilter type="%s"><float name="%s" value="%s"/></rfilter>' % (t, n, v), txt, count=1)
        if k != 1:
            raise RuntimeError("no <rfilter .../> to replace in " + xml_path)
    if subs.get("extra_xml"):
        txt = txt.replace("</scene>", subs["extra_xml"] + "\n</scene>")
    out = os.path.join(out_dir, subs.get("name", "variant") + ".xml")
    with open(out, "w") as f:
        f.write(txt)
    return out


class RefScene:
    def __init__(self, xml_path, threads=None):
        init(threads)
        with _quiet():
            self.h = lib().gdpt_ref_load(os.path.abspath(xml_path).encode())
        if not self.h:
            raise RuntimeError("reference failed to parse " + xml_path)
        self.refresh()

    def refresh(self):
        info = (C.c_int * 8)()
        lib().gdpt_ref_info(self.h, info)
        (self.width, self.height, self.spp, self.max_depth, self.rr_depth,
         self.integrator, self.nshapes, self.nmats) = list(info)

    def set_options(self, spp=-2, max_depth=-2, rr_depth=-2):
        lib().gdpt_ref_set_options(self.h, spp, max_depth, rr_depth)
        self.refresh()

    def close(self):
        if self.h:
            lib().gdpt_ref_free(self.h)
            self.h = None

    def _img(self):
        return np.zeros((self.height, self.width, 3), dtype=np.float64)

    def render(self):
        """Reference render() with GradPath. Returns dict(primal, dx, dy, poisson) = c, cx, cy, out."""
        c, cx, cy, out = self._img(), self._img(), self._img(), self._img()
        with _quiet():
            lib().gdpt_ref_render(self.h, _dp(c), _dp(cx), _dp(cy), _dp(out))
        return dict(primal=c, dx=cx, dy=cy, poisson=out)

    def render_path(self):
        out = self._img()
        with _quiet():
            lib().gdpt_ref_render_path(self.h, _dp(out))
        return out

    def render_aux(self, integrator):
        """aux_render (src/render.cpp:15-72) for Integrator value 0..4."""
        out = self._img()
        with _quiet():
            lib().gdpt_ref_render_aux(self.h, int(integrator), _dp(out))
        return out

    def grad_sample(self, x, y, stream, seed=DEFAULT_SEED):
        rec = np.zeros(23)
        lib().gdpt_ref_grad_sample(self.h, x, y, stream, seed, _dp(rec))
        return rec

    def render_pixelseed(self, spp, seed=DEFAULT_SEED, sample0=0, y0=0, y1=None):
        """Pixel-seeded accumulation (stream = sample*W*H + y*W + x). Returns c, cx0, cy0, cx1, cy1."""
        y1 = self.height if y1 is None else y1
        bufs = [self._img() for _ in range(5)]
        lib().gdpt_ref_render_pixelseed(self.h, seed, spp, sample0, y0, y1, *[_dp(b) for b in bufs])
        return bufs

    def sample_primary(self, screen_pos):
        sp = np.ascontiguousarray(screen_pos, dtype=np.float64).reshape(-1, 2)
        rays = np.zeros((sp.shape[0], 6))
        lib().gdpt_ref_sample_primary(self.h, sp.shape[0], _dp(sp), _dp(rays))
        return rays

    def intersect(self, rays, rdiff=None):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 8)
        n = rays.shape[0]
        ids = np.zeros((n, 3), dtype=np.int32)
        vtx = np.zeros((n, 24))
        rd = np.ascontiguousarray(rdiff, dtype=np.float64) if rdiff is not None else None
        lib().gdpt_ref_intersect(self.h, n, _dp(rays), _dp(rd), ids.ctypes.data_as(C.POINTER(C.c_int)), _dp(vtx))
        return ids, vtx

    def bsdf(self, mat_id, vtx, dirs, rnd):
        vtx = np.ascontiguousarray(vtx, dtype=np.float64).reshape(-1, 15)
        dirs = np.ascontiguousarray(dirs, dtype=np.float64).reshape(-1, 6)
        rnd = np.ascontiguousarray(rnd, dtype=np.float64).reshape(-1, 3)
        out = np.zeros((vtx.shape[0], 10))
        lib().gdpt_ref_bsdf(self.h, mat_id, vtx.shape[0], _dp(vtx), _dp(dirs), _dp(rnd), _dp(out))
        return out

    def path_sample(self, x, y, stream, seed=DEFAULT_SEED):
        """One path_tracing evaluation (src/path_tracing.h:13-348) with rng = init_pcg32(stream, seed)."""
        rad = np.zeros(3)
        lib().gdpt_ref_path_sample(self.h, x, y, stream, seed, _dp(rad))
        return rad

    def shift_sample(self, x, y, stream, seed=DEFAULT_SEED):
        """One shift-mapped gradpath sample (the product's opt-in offset mode) restated on the reference's own primitives
        (ref_hooks.cpp: gdpt_ref_shift_sample) -> the 23-double record, offsets X0=(x-1,y), Y0=(x,y-1), X1, Y1."""
        rec = np.zeros(23)
        lib().gdpt_ref_shift_sample(self.h, x, y, stream, seed, _dp(rec))
        return rec

    def render_path_pixelseed(self, spp, seed=DEFAULT_SEED, sample0=0):
        img = self._img()
        lib().gdpt_ref_render_path_pixelseed(self.h, seed, spp, sample0, _dp(img))
        return img

    def sample_light(self, ref_points, rnd):
        ref_points = np.ascontiguousarray(ref_points, dtype=np.float64).reshape(-1, 3)
        rnd = np.ascontiguousarray(rnd, dtype=np.float64).reshape(-1, 4)
        out = np.zeros((ref_points.shape[0], 9))
        lib().gdpt_ref_sample_light(self.h, ref_points.shape[0], _dp(ref_points), _dp(rnd), _dp(out))
        return out

    def occluded(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 8)
        out = np.zeros(rays.shape[0], dtype=np.int32)
        lib().gdpt_ref_occluded(self.h, rays.shape[0], _dp(rays), out.ctypes.data_as(C.POINTER(C.c_int)))
        return out.astype(bool)

    def export(self, path):
        if lib().gdpt_ref_export(self.h, os.path.abspath(path).encode()) != 0:
            raise RuntimeError("export failed: " + path)
        return path


def imread3(path):
    """The reference's own image loader (src/image.cpp:80-133) on a file."""
    w, h = C.c_int(0), C.c_int(0)
    p = os.path.abspath(path).encode()
    lib().gdpt_ref_imread3(p, C.byref(w), C.byref(h), None)
    img = np.zeros((h.value, w.value, 3))
    lib().gdpt_ref_imread3(p, C.byref(w), C.byref(h), _dp(img))
    return img


def fourier_solve(u, gx, gy, alpha=0.04):
    u = np.ascontiguousarray(u, dtype=np.float64)
    gx = np.ascontiguousarray(gx, dtype=np.float64)
    gy = np.ascontiguousarray(gy, dtype=np.float64)
    h, w, _ = u.shape
    out = np.zeros_like(u)
    lib().gdpt_ref_fourier_solve(w, h, _dp(u), _dp(gx), _dp(gy), alpha, _dp(out))
    return out


# ---------------------------------------------------------------------------------------
# Staged scenes: copies of the reference's scene ASSETS (XML/OBJ/.serialized/JPG — data,
# not source) plus generated variant XMLs, under oracle/_ref/scenes/ (git-ignored, but it
# travels to the GPU box with oracle/_ref/).  Only test infrastructure reads them.
STAGE_DIR = os.path.join(_HERE, "_ref", "scenes")

AREA_EMITTER_XML = """
	<shape type="sphere">
		<point name="center" x="1.5" y="-1.5" z="6"/>
		<float name="radius" value="1.0"/>
		<emitter type="area">
			<rgb name="radiance" value="40, 40, 40"/>
		</emitter>
		<bsdf type="diffuse">
			<rgb name="reflectance" value="0,0,0"/>
		</bsdf>
	</shape>
"""

# name -> (scene dir, source xml, substitutions)
VARIANTS = {
    "cbox_gdpt": ("cbox", "cbox_gdpt.xml", dict()),
    "cbox_gdpt_64": ("cbox", "cbox_gdpt.xml", dict(width=64, height=64)),
    "cbox_gdpt_d2": ("cbox", "cbox_gdpt.xml", dict(width=64, height=64, max_depth=2)),
    "veach_mi_gradpath": ("veach_mi", "mi.xml", dict(integrator="gradpath", max_depth=-1, spp=64)),
    "veach_mi_small": ("veach_mi", "mi.xml", dict(integrator="gradpath", max_depth=-1, spp=4, width=96, height=64)),
    "disney_gradpath": ("disney_bsdf_test", "disney_bsdf.xml", dict(integrator="gradpath", spp=256, extra_xml=AREA_EMITTER_XML)),
    "disney_small": ("disney_bsdf_test", "disney_bsdf.xml", dict(integrator="gradpath", spp=4, width=96, height=72, extra_xml=AREA_EMITTER_XML)),
    "matpreview_small": ("matpreview", "matpreview.xml", dict(width=96, height=72, spp=4)),
    "sponza_1080p": ("sponza", "sponza.xml", dict(integrator="gradpath", width=1920, height=1080, spp=512)),
    "sponza_2160p": ("sponza", "sponza.xml", dict(integrator="gradpath", width=3840, height=2160, spp=1024)),
    "sponza_native": ("sponza", "sponza.xml", dict(integrator="gradpath")),
    "sponza_small": ("sponza", "sponza.xml", dict(integrator="gradpath", width=160, height=90, spp=4)),
    "cbox_gdpt_64_box": ("cbox", "cbox_gdpt.xml", dict(width=64, height=64, rfilter=("box", "width", "2"))),
    "cbox_gdpt_64_tent": ("cbox", "cbox_gdpt.xml", dict(width=64, height=64, rfilter=("tent", "width", "2"))),
    "cbox_gdpt_64_gauss2": ("cbox", "cbox_gdpt.xml", dict(width=64, height=64, rfilter=("gaussian", "stddev", "2"))),
}


def stage_scenes():
    """Copy the scene assets next to generated variant XMLs. Needs /root/reference."""
    import shutil
    src_root = os.path.join(REFERENCE_DIR, "scenes")
    if not os.path.isdir(src_root):
        return False
    for d in ("cbox", "veach_mi", "sponza", "matpreview", "disney_bsdf_test"):
        dst = os.path.join(STAGE_DIR, d)
        if not os.path.isdir(dst):
            shutil.copytree(os.path.join(src_root, d), dst, ignore=shutil.ignore_patterns(".DS_Store"))
    for name, (d, xml, subs) in VARIANTS.items():
        make_variant(os.path.join(STAGE_DIR, d, xml), out_dir=os.path.join(STAGE_DIR, d), name="gdpt_" + name, **subs)
    return True


def scene_path(name):
    d, _, _ = VARIANTS[name]
    p = os.path.join(STAGE_DIR, d, "gdpt_" + name + ".xml")
    if not os.path.exists(p):
        if not stage_scenes():
            raise RuntimeError("staged scene %s missing and /root/reference not available" % name)
    return p


def scenes_available():
    return available() and (os.path.isdir(STAGE_DIR) or os.path.isdir(os.path.join(REFERENCE_DIR, "scenes")))
