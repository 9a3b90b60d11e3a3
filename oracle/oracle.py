"""TEST INFRASTRUCTURE ONLY — never imported by the product package.

ctypes front-end for the two CPU checkers that implement oracle/oracle_api.h:
  Oracle("port")  -> oracle/libbdpt_port.so        (our restatement, oracle/port.cpp)
  Oracle("ref")   -> oracle/_ref/libref_oracle.so  (the reference's own headers + oracle/ref_harness.cpp)
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_LIGHT = 0, 1, 2, 3
RNG_MT19937, RNG_PHILOX = 0, 1
WEIGHT_REF_SUM, WEIGHT_MIS = 0, 1
MISS = 0xFFFFFFFF


class Material(C.Structure):
    _fields_ = [("type", C.c_uint32), ("tex", C.c_uint32), ("rgb", C.c_double * 3), ("param", C.c_double)]


class Texture(C.Structure):
    _fields_ = [("type", C.c_uint32), ("pad", C.c_uint32), ("a", C.c_double * 3), ("b", C.c_double * 3), ("scale", C.c_double),
                ("data_offset", C.c_uint64), ("width", C.c_int32), ("height", C.c_int32)]


TEX_SOLID, TEX_CHECKER, TEX_NOISE, TEX_IMAGE = 0, 1, 2, 3
NOISE_BYTES = 256 * 3 * 8 + 3 * 256 * 4


class Camera(C.Structure):
    _fields_ = [("aspect_ratio", C.c_double), ("image_width", C.c_int32), ("samples_per_pixel", C.c_int32),
                ("max_depth", C.c_int32), ("pad", C.c_int32), ("background", C.c_double * 3),
                ("vfov", C.c_double), ("lookfrom", C.c_double * 3), ("lookat", C.c_double * 3),
                ("vup", C.c_double * 3), ("defocus_angle", C.c_double), ("focus_dist", C.c_double)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays", "node_visits", "aabb_tests", "aabb_hits", "tri_tests",
                                          "tri_hits", "nodes_built", "shadow_rays")]


def make_camera(width, height, spp, depth, lookfrom, lookat, vfov=90.0, vup=(0, 1, 0), background=(0, 0, 0),
                focus_dist=10.0):
    c = Camera()
    c.aspect_ratio = float(width) / float(height)  # scene_loader.h:445
    while max(1, int(width / c.aspect_ratio)) != int(height):   # camera.h:161 truncates: ~2 % of (W, H) pairs lose a row
        c.aspect_ratio = float(np.nextafter(c.aspect_ratio, 0.0))
    c.image_width, c.samples_per_pixel, c.max_depth = int(width), int(spp), int(depth)
    c.background[:] = background
    c.vfov = vfov
    c.lookfrom[:] = lookfrom
    c.lookat[:] = lookat
    c.vup[:] = vup
    c.defocus_angle = 0.0
    c.focus_dist = focus_dist
    return c


def make_materials(mats):
    """mats: list of (type, (r,g,b), param[, tex]) -- tex = 1-based index into the texture table, 0 / absent = none"""
    arr = (Material * len(mats))()
    for i, mt in enumerate(mats):
        arr[i].type = mt[0]
        arr[i].rgb[:] = mt[1]
        arr[i].param = mt[2]
        arr[i].tex = mt[3] if len(mt) > 3 else 0
    return arr


def make_textures(textures):
    """textures: list of dicts {type, a, b, scale, width, height, data (bytes / uint8 array / None)} -> (Texture array, blob).
    Noise textures without data get room for their tables (filled later from the reference: Oracle.texture_data)."""
    arr = (Texture * max(1, len(textures)))()
    blob = bytearray()
    for i, t in enumerate(textures):
        arr[i].type = t["type"]
        arr[i].a[:] = t.get("a", (0, 0, 0))
        arr[i].b[:] = t.get("b", (0, 0, 0))
        arr[i].scale = t.get("scale", 1.0)
        arr[i].width, arr[i].height = t.get("width", 0), t.get("height", 0)
        while len(blob) % 8:
            blob.append(0)
        arr[i].data_offset = len(blob)
        d = t.get("data")
        if d is not None:
            blob += bytes(np.ascontiguousarray(d).tobytes() if not isinstance(d, (bytes, bytearray)) else d)
        elif t["type"] == TEX_NOISE:
            blob += bytes(NOISE_BYTES)
    return arr, np.frombuffer(bytes(blob) + b"\0" * 8, dtype=np.uint8).copy()


def build(which="all"):
    subprocess.check_call(["make", "-C", HERE, "-s"] + ([] if which == "all" else [which]))


def lib_path(kind):
    return os.path.join(HERE, "libbdpt_port.so" if kind == "port" else os.path.join("_ref", "libref_oracle.so"))


def available(kind):
    return os.path.exists(lib_path(kind))


class Oracle:
    def __init__(self, kind="port"):
        self.kind = kind
        path = lib_path(kind)
        if not os.path.exists(path):
            build("libbdpt_port.so" if kind == "port" else "ref")
        self.lib = C.CDLL(path)
        p = "port_" if kind == "port" else "ref_"
        f = lambda n: getattr(self.lib, p + n)
        self._create = f("scene_create"); self._create.restype = C.c_void_p
        self._create.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32]
        self._create_tex = f("scene_create_tex"); self._create_tex.restype = C.c_void_p
        self._create_tex.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        self._destroy = f("scene_destroy"); self._destroy.argtypes = [C.c_void_p]
        self._nodes = f("bvh_node_count"); self._nodes.restype = C.c_uint64; self._nodes.argtypes = [C.c_void_p]
        self._order = f("bvh_leaf_order"); self._order.argtypes = [C.c_void_p, C.c_void_p]
        self._trace = f("trace"); self._trace.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64] + [C.c_void_p] * 4 + [C.c_int]
        self._visible = f("visible"); self._visible.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
        self._render = f("render"); self._render.restype = C.c_int
        self._render.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_int]
        if kind == "port":
            self._render_ex = self.lib.port_render_ex; self._render_ex.restype = C.c_int
            self._render_ex.argtypes = self._render.argtypes + [C.c_int]
        else:
            self._render_pt = self.lib.ref_render_pt; self._render_pt.restype = C.c_int
            self._render_pt.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int, C.c_void_p]
        self._stats = f("stats_get"); self._stats.argtypes = [C.c_void_p]
        self._stats_reset = f("stats_reset")
        self._rgb8 = f("colors_to_rgb8"); self._rgb8.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        self.scene = None

    def load(self, verts, mat_id, mats, textures=None):
        """textures: (Texture array, n, uint8 blob) as make_textures returns them with the count, or None."""
        verts = np.ascontiguousarray(verts, dtype=np.float64).reshape(-1, 9)
        mat_id = np.ascontiguousarray(mat_id, dtype=np.uint32)
        assert len(mat_id) == len(verts)
        self.n_tris = len(verts)
        marr = make_materials(mats) if isinstance(mats, list) else mats
        if self.scene:
            self._destroy(self.scene)
        if textures is not None:
            tarr, n_tex, blob = textures
            self.scene = self._create_tex(verts.ctypes.data, mat_id.ctypes.data, len(verts), C.byref(marr), len(marr), C.byref(tarr), n_tex,
                                          blob.ctypes.data, len(blob))
        else:
            self.scene = self._create(verts.ctypes.data, mat_id.ctypes.data, len(verts), C.byref(marr), len(marr))
        return self

    def set_media(self, media):
        """media: list of (density, (r, g, b), boundary triangles as n x 9) -- constant_medium + isotropic."""
        rows, soup, first = [], [], 0
        for density, rgb, tris in media:
            t = np.ascontiguousarray(tris, dtype=np.float64).reshape(-1, 9)
            rows.append([float(density), float(rgb[0]), float(rgb[1]), float(rgb[2]), float(first), float(len(t))])
            soup.append(t); first += len(t)
        m6 = np.ascontiguousarray(rows, dtype=np.float64).reshape(-1, 6)
        b = np.ascontiguousarray(np.concatenate(soup) if soup else np.zeros((0, 9)), dtype=np.float64)
        f = getattr(self.lib, ("port_" if self.kind == "port" else "ref_") + "scene_set_media")
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
        f.restype = None
        f(self.scene, m6.ctypes.data, len(m6), b.ctypes.data)
        return self

    def set_spheres(self, spheres):
        """port only.  spheres: list of ((cx, cy, cz), radius, material index) -- analytic sphere prims (an extension)."""
        assert self.kind == "port"
        a = np.ascontiguousarray([[c[0], c[1], c[2], r, m] for c, r, m in spheres], dtype=np.float64).reshape(-1, 5)
        f = self.lib.port_scene_set_spheres
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        f.restype = None
        f(self.scene, a.ctypes.data, len(a))
        return self

    def texture_value(self, k, uvp):
        uvp = np.ascontiguousarray(uvp, dtype=np.float64).reshape(-1, 5)
        out = np.empty((len(uvp), 3), np.float64)
        f = getattr(self.lib, ("port_" if self.kind == "port" else "ref_") + "texture_value")
        f.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p, C.c_void_p]
        f.restype = None
        f(self.scene, k, len(uvp), uvp.ctypes.data, out.ctypes.data)
        return out

    def texture_data(self, k, nbytes):
        """ref only: what the reference's own texture object k holds (perlin tables / image bytes after its float round trip)."""
        assert self.kind == "ref"
        out = np.zeros(nbytes, np.uint8)
        f = self.lib.ref_texture_data
        f.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        rc = f(self.scene, k, out.ctypes.data)
        assert rc == 0, rc
        return out

    def node_count(self):
        return int(self._nodes(self.scene))

    def leaf_order(self):
        out = np.empty(self.n_tris, dtype=np.uint32)
        self._order(self.scene, out.ctypes.data)
        return out

    def trace(self, rays, nthreads=1):
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 8)
        n = len(rays)
        prim = np.empty(n, np.uint32); t = np.empty(n); u = np.empty(n); v = np.empty(n)
        self._trace(self.scene, rays.ctypes.data, n, prim.ctypes.data, t.ctypes.data, u.ctypes.data, v.ctypes.data, nthreads)
        return prim, t, u, v

    def visible(self, a, b, nthreads=1):
        a = np.ascontiguousarray(a, dtype=np.float64).reshape(-1, 3)
        b = np.ascontiguousarray(b, dtype=np.float64).reshape(-1, 3)
        out = np.empty(len(a), np.uint8)
        self._visible(self.scene, a.ctypes.data, b.ctypes.data, len(a), out.ctypes.data, nthreads)
        return out.astype(bool)

    def render(self, cam, rng_mode=RNG_MT19937, seed=0, row0=0, row1=1 << 30, nthreads=1, height=None, weighting=WEIGHT_REF_SUM):
        W = cam.image_width
        H = height if height is not None else max(1, int(W / cam.aspect_ratio))
        film = np.zeros((H, W, 3), dtype=np.float64)
        if weighting != WEIGHT_REF_SUM:
            assert self.kind == "port", "the reference has one estimator; MIS exists in the port only"
            h = self._render_ex(self.scene, C.byref(cam), rng_mode, seed, row0, row1, film.ctypes.data, nthreads, weighting)
        else:
            h = self._render(self.scene, C.byref(cam), rng_mode, seed, row0, row1, film.ctypes.data, nthreads)
        assert h == H, (h, H)
        return film

    def render_pt(self, cam, seed=0, row0=0, row1=1 << 30, height=None):
        """ref only: the reference's own unbiased path tracer (camera::path_trace_color, camera.h:255-292)."""
        assert self.kind == "ref"
        W = cam.image_width
        H = height if height is not None else max(1, int(W / cam.aspect_ratio))
        film = np.zeros((H, W, 3), dtype=np.float64)
        h = self._render_pt(self.scene, C.byref(cam), seed, row0, row1, film.ctypes.data)
        assert h == H, (h, H)
        return film

    def stats(self):
        s = Stats()
        self._stats(C.byref(s))
        return {n: int(getattr(s, n)) for n, _ in Stats._fields_}

    def stats_reset(self):
        self._stats_reset()

    def colors_to_rgb8(self, film, spp):
        film = np.ascontiguousarray(film, dtype=np.float64)
        H, W, _ = film.shape
        out = np.empty((H, W, 3), np.uint8)
        self._rgb8(film.ctypes.data, W, H, spp, out.ctypes.data)
        return out

    def close(self):
        if self.scene:
            self._destroy(self.scene)
            self.scene = None
