"""Host-side scene model: the POD mirror of lajolla's `Scene` (src/scene.h:43-83) that the
C ABI consumes, loaded from a "GDPT1" array container (written by the reference's own
parser through oracle/ref.py::RefScene.export, or by tools/).

Container: 8-byte magic "GDPT1\\0\\0\\0", then records
  u32 name_len, name, char dtype[8], u32 ndim, i64 shape[ndim], u64 nbytes, payload.
A file ending in ".z" is the same stream zlib-compressed.
"""
import ctypes as C
import enum
import struct
import zlib

import numpy as np

from . import _capi


class Integrator(enum.IntEnum):
    """Same enumerators, same order, as lajolla's `enum class Integrator` (src/scene.h:14-23)."""
    Depth = 0
    ShadingNormal = 1
    MeanCurvature = 2
    RayDifferential = 3
    MipmapLevel = 4
    Path = 5
    VolPath = 6
    GradPath = 7


_DTYPES = {"f8": np.float64, "f4": np.float32, "i4": np.int32, "i8": np.int64, "u1": np.uint8, "u2": np.uint16}


def read_container(path):
    raw = open(path, "rb").read()
    if path.endswith(".z"):
        raw = zlib.decompress(raw)
    if raw[:5] != b"GDPT1":
        raise ValueError("%s is not a GDPT1 container" % path)
    pos, out = 8, {}
    while pos < len(raw):
        (nl,) = struct.unpack_from("<I", raw, pos); pos += 4
        name = raw[pos:pos + nl].decode(); pos += nl
        dt = raw[pos:pos + 8].split(b"\0")[0].decode(); pos += 8
        (nd,) = struct.unpack_from("<I", raw, pos); pos += 4
        shape = struct.unpack_from("<%dq" % nd, raw, pos); pos += 8 * nd
        (nb,) = struct.unpack_from("<Q", raw, pos); pos += 8
        out[name] = np.frombuffer(raw, dtype=_DTYPES[dt], count=nb // np.dtype(_DTYPES[dt]).itemsize, offset=pos).reshape(shape).copy()
        pos += nb
    return out


def write_container(path, arrays):
    rev = {np.dtype(v): k for k, v in _DTYPES.items()}
    chunks = [b"GDPT1\0\0\0"]
    for name, a in arrays.items():
        a = np.ascontiguousarray(a)
        dt = rev[a.dtype].encode().ljust(8, b"\0")
        chunks.append(struct.pack("<I", len(name)) + name.encode() + dt + struct.pack("<I", a.ndim) +
                      struct.pack("<%dq" % a.ndim, *a.shape) + struct.pack("<Q", a.nbytes) + a.tobytes())
    raw = b"".join(chunks)
    if path.endswith(".z"):
        raw = zlib.compress(raw, 9)
    with open(path, "wb") as f:
        f.write(raw)


# camera/options-only containers and the full container they sit on
OVERLAY_BASE = {"sponza_2160p": "sponza_1080p", "sponza_native": "sponza_1080p", "sponza_small": "sponza_1080p",
                "disney_small": "disney_gradpath",
                # pixel-filter variants of the 64x64 Cornell box: Box(width 2), Tent(width 2), Gaussian(stddev 2) — the three
                # filters of src/tests/filter.cpp:15-66
                "cbox_gdpt_64_box": "cbox_gdpt_64", "cbox_gdpt_64_tent": "cbox_gdpt_64", "cbox_gdpt_64_gauss2": "cbox_gdpt_64"}


class RenderOptions:
    """src/scene.h:25-32"""

    def __init__(self, integrator=Integrator.GradPath, samples_per_pixel=4, max_depth=-1, rr_depth=5):
        self.integrator = Integrator(integrator)
        self.samples_per_pixel = samples_per_pixel
        self.max_depth = max_depth
        self.rr_depth = rr_depth


class Scene:
    """Flattened scene + the device context built from it (lazily, per device)."""

    def __init__(self, arrays):
        self.a = arrays
        w, h = (int(v) for v in arrays["camera.size"])
        self.width, self.height = w, h
        o = arrays["options"]
        self.options = RenderOptions(Integrator(int(o[3])), int(o[0]), int(o[1]), int(o[2]))
        self._ctx = {}
        self._keep = []

    @classmethod
    def load(cls, path, overlay=None):
        """overlay: a second container whose arrays (camera.*, options) replace the first's —
        e.g. scenes/sponza_2160p.camera.gdpt on top of scenes/sponza_1080p.gdpt.z."""
        arrays = read_container(path)
        if overlay:
            arrays.update(read_container(overlay))
        return cls(arrays)

    @classmethod
    def named(cls, name):
        """Load one of the containers shipped under scenes/ by config name."""
        import os
        root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes")
        full = os.path.join(root, name + ".gdpt.z")
        if os.path.exists(full):
            return cls.load(full)
        cam = os.path.join(root, name + ".camera.gdpt")
        if os.path.exists(cam) and name in OVERLAY_BASE:
            return cls.load(os.path.join(root, OVERLAY_BASE[name] + ".gdpt.z"), overlay=cam)
        raise FileNotFoundError("no scene container named " + name)

    def n_images(self, prefix):
        n = 0
        while "%s.%d.size" % (prefix, n) in self.a:
            n += 1
        return n

    def _images(self, prefix, channels):
        n = self.n_images(prefix)
        arr = (_capi.GdptImage * max(n, 1))()
        for i in range(n):
            w, h = (int(v) for v in self.a["%s.%d.size" % (prefix, i)])
            g = arr[i]
            g.width, g.height, g.channels = w, h, channels
            if "%s.%d.data" % (prefix, i) in self.a:
                d = np.ascontiguousarray(self.a["%s.%d.data" % (prefix, i)], dtype=np.float64)
                self._keep.append(d)
                g.data = _capi.as_dp(d)
            else:
                pal = np.ascontiguousarray(self.a["%s.%d.palette" % (prefix, i)], dtype=np.float64)
                idx = np.ascontiguousarray(self.a["%s.%d.index" % (prefix, i)], dtype=np.uint8)
                self._keep += [pal, idx]
                g.palette = _capi.as_dp(pal)
                g.palette_size = pal.size
                g.index = idx.ctypes.data_as(C.POINTER(C.c_uint8))
        return n, arr

    def desc(self):
        a, d = self.a, _capi.GdptSceneDesc()
        keep = self._keep

        def dbl(name, cols):
            v = np.ascontiguousarray(a[name], dtype=np.float64).reshape(-1, cols) if a[name].size else np.zeros((1, cols))
            keep.append(v)
            return v

        d.sample_to_cam[:] = list(np.asarray(a["camera.sample_to_cam"], dtype=np.float64).ravel())
        d.cam_to_world[:] = list(np.asarray(a["camera.cam_to_world"], dtype=np.float64).ravel())
        d.width, d.height = self.width, self.height
        d.filter_kind, d.filter_param = int(a["camera.filter"][0]), float(a["camera.filter"][1])
        shapes = np.ascontiguousarray(a["shapes"], dtype=np.int32); keep.append(shapes)
        offs = np.ascontiguousarray(a["shape_offsets"], dtype=np.int64); keep.append(offs)
        d.n_shapes = shapes.shape[0]
        d.shapes = shapes.ctypes.data_as(_capi.ip)
        d.shape_offsets = offs.ctypes.data_as(C.POINTER(C.c_int64))
        d.spheres = _capi.as_dp(dbl("spheres", 4))
        d.n_positions, d.n_triangles = a["positions"].shape[0], a["indices"].shape[0]
        d.n_normals, d.n_uvs = a["normals"].shape[0], a["uvs"].shape[0]
        d.positions = _capi.as_dp(dbl("positions", 3))
        idx = np.ascontiguousarray(a["indices"], dtype=np.int32) if a["indices"].size else np.zeros((1, 3), np.int32)
        keep.append(idx)
        d.indices = idx.ctypes.data_as(_capi.ip)
        d.normals = _capi.as_dp(dbl("normals", 3))
        d.uvs = _capi.as_dp(dbl("uvs", 2))
        d.n_materials = a["materials"].shape[0]
        d.materials = _capi.as_dp(dbl("materials", 14))
        d.n_textures = a["textures"].shape[0]
        d.textures = _capi.as_dp(dbl("textures", 14))
        d.n_images3, img3 = self._images("img3", 3)
        d.n_images1, img1 = self._images("img1", 1)
        keep += [img3, img1]
        d.images3, d.images1 = img3, img1
        d.n_lights = a["lights"].shape[0]
        d.lights = _capi.as_dp(dbl("lights", 5))
        d.max_depth, d.rr_depth = self.options.max_depth, self.options.rr_depth
        d.samples_per_pixel = self.options.samples_per_pixel
        if "envmap.to_world" in a and "envmap.to_local" in a:      # Envmap::to_world / to_local (src/light.h:23-30)
            d.envmap_to_world = _capi.as_dp(dbl("envmap.to_world", 4))
            d.envmap_to_local = _capi.as_dp(dbl("envmap.to_local", 4))
        return d

    def context(self, device=0, build_bvh=True, bvh=None):
        """gdpt_create on `device` (once per scene and device).  build_bvh=False: gdpt_create_ex(GDPT_CREATE_NO_BVH) — the
        caller hands a tree over with adopt_bvh() before rendering (multi.DistributedRenderer broadcasts rank 0's)."""
        if device not in self._ctx:
            h = C.c_void_p()
            d = self.desc()
            flags = 0 if build_bvh else _capi.CREATE_NO_BVH      # bvh: None (chosen from the planned work) | "host" | "device"
            flags |= {None: 0, "host": _capi.CREATE_BVH_HOST, "device": _capi.CREATE_BVH_DEVICE}[bvh]
            _capi.check(_capi.lib().gdpt_create_ex(C.byref(d), device, flags, C.byref(h)))
            self._ctx[device] = h
        return self._ctx[device]

    def bvh_info(self, device=0):
        """(n_nodes, n_prims, depth, device pointer of the nodes, device pointer of the primitive records)"""
        nn, npr, dep, pn, pp = C.c_int64(), C.c_int64(), C.c_int32(), C.c_void_p(), C.c_void_p()
        _capi.check(_capi.lib().gdpt_bvh_info(self.context(device), C.byref(nn), C.byref(npr), C.byref(dep), C.byref(pn), C.byref(pp)))
        return nn.value, npr.value, dep.value, pn.value, pp.value

    def adopt_bvh(self, nodes_ptr, n_nodes, prims_ptr, n_prims, depth, device=0):
        _capi.check(_capi.lib().gdpt_bvh_adopt(self._ctx[device], C.c_void_p(nodes_ptr), n_nodes, C.c_void_p(prims_ptr), n_prims, depth))

    def set_option(self, name, value, device=0):
        """gdpt_set_option: scheduling / tuning knobs of the context ("async_waves", "wave_streams", "time_kernels", ...)."""
        _capi.check(_capi.lib().gdpt_set_option(self.context(device), name.encode(), int(value)))

    def close(self):
        for h in self._ctx.values():
            _capi.lib().gdpt_destroy(h)
        self._ctx = {}

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
