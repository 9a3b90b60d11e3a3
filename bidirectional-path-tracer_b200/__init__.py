"""B200-native BDPT hot path — thin ctypes binding over the C-ABI of include/bdpt.h.

The product is libbdpt.so (hand-written CUDA for sm_100a + host C++); this module only marshals numpy
buffers across the C boundary for tests and bench.py.  There is no CPU fallback: creating a `Context`
without a CUDA device raises.  The directory name carries a hyphen, so import it with
`importlib.import_module("bidirectional-path-tracer_b200")` (see __graft_entry__.load_package()).
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BDPT_LIB") or os.path.join(HERE, "libbdpt.so")   # BDPT_LIB: a differently built library (development A/B runs)

MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_LIGHT = 0, 1, 2, 3
BVH_REFERENCE, BVH_FAST = 0, 1
WEIGHT_REF_SUM, WEIGHT_MIS = 0, 1
TRACE_CLOSEST, TRACE_ANY, TRACE_DEVICE_PTRS, TRACE_STATS, TRACE_SPLIT = 0, 1, 16, 32, 64
MISS = 0xFFFFFFFF
NOALLOC_RGB8_MAX_BYTES = 32 << 20   # BDPT_NOALLOC_RGB8_MAX_BYTES

RAY_DTYPE = np.dtype([("o", "<f8", 3), ("d", "<f8", 3), ("tmin", "<f8"), ("tmax", "<f8")])
HIT_DTYPE = np.dtype([("t", "<f8"), ("u", "<f8"), ("v", "<f8"), ("prim", "<u4"), ("pad", "<u4")])

EXPORTS = [
    "bdpt_ctx_create", "bdpt_ctx_destroy", "bdpt_last_error", "bdpt_scene_upload", "bdpt_scene_upload_ex", "bdpt_bvh_info",
    "bdpt_bvh_leaf_order", "bdpt_host_bvh_build", "bdpt_trace", "bdpt_visible", "bdpt_render",
    "bdpt_film_clear", "bdpt_film_device_ptr", "bdpt_film_download", "bdpt_film_size", "bdpt_film_to_rgb8", "bdpt_stats",
    "bdpt_stats_reset", "bdpt_last_device_ms", "bdpt_sync", "bdpt_kernel_times",
    "bdpt_scene_file_load", "bdpt_scene_file_free", "bdpt_scene_file_info", "bdpt_scene_file_copy",
    "bdpt_write_png", "bdpt_host_last_error",
    "bdpt_scene_image_build", "bdpt_scene_image_free", "bdpt_scene_image_info", "bdpt_scene_image_last_error",
    "bdpt_scene_upload_image", "bdpt_scene_image_save", "bdpt_scene_image_map", "bdpt_scene_image_unmap",
    "bdpt_scene_source_key", "bdpt_scene_file_open_cached", "bdpt_test_filter32",
    "bdpt_scene_set_textures", "bdpt_scene_file_textures", "bdpt_film_reduce", "bdpt_comm_unique_id", "bdpt_comm_init", "bdpt_film_reduce_comm", "bdpt_render_progress", "bdpt_bench_read", "bdpt_comm_attach",
    "bdpt_film_to_rgb8_noalloc", "bdpt_scene_set_media", "bdpt_scene_file_media", "bdpt_scene_set_spheres", "bdpt_scene_file_spheres",
]


class Material(C.Structure):
    _fields_ = [("type", C.c_uint32), ("tex", C.c_uint32), ("rgb", C.c_double * 3), ("param", C.c_double)]


class Sphere(C.Structure):
    _fields_ = [("center", C.c_double * 3), ("radius", C.c_double), ("mat_id", C.c_uint32), ("pad", C.c_uint32)]


class Medium(C.Structure):
    _fields_ = [("density", C.c_double), ("albedo", C.c_double * 3), ("first_tri", C.c_uint64), ("n_tris", C.c_uint64)]


class Texture(C.Structure):
    _fields_ = [("type", C.c_uint32), ("pad", C.c_uint32), ("a", C.c_double * 3), ("b", C.c_double * 3), ("scale", C.c_double),
                ("data_offset", C.c_uint64), ("width", C.c_int32), ("height", C.c_int32)]


TEX_SOLID, TEX_CHECKER, TEX_NOISE, TEX_IMAGE = 0, 1, 2, 3


class Camera(C.Structure):
    _fields_ = [("aspect_ratio", C.c_double), ("image_width", C.c_int32), ("samples_per_pixel", C.c_int32),
                ("max_depth", C.c_int32), ("pad", C.c_int32), ("background", C.c_double * 3),
                ("vfov", C.c_double), ("lookfrom", C.c_double * 3), ("lookat", C.c_double * 3),
                ("vup", C.c_double * 3), ("defocus_angle", C.c_double), ("focus_dist", C.c_double)]


class RenderOpts(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("stratum_begin", C.c_int32), ("stratum_end", C.c_int32),
                ("stratum_step", C.c_int32), ("bvh", C.c_int32), ("row_begin", C.c_int32),
                ("row_end", C.c_int32), ("batch_samples", C.c_int64), ("flags", C.c_int32), ("streams", C.c_int32),
                ("weighting", C.c_int32), ("pad", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("rays_closest", "rays_shadow", "node_visits", "tri_tests", "samples",
                                          "kernel_launches", "camera_vertices", "light_vertices", "nodes_closest",
                                          "tris_closest", "nodes_shadow", "tris_shadow", "max_work_closest",
                                          "max_work_shadow", "deferred_closest", "deferred_shadow", "aabb_tests", "aabb_hits",
                                          "triangle_hits", "bvh_nodes_built")]


class SceneImageInfo(C.Structure):
    _fields_ = [("n_tris", C.c_uint64), ("n_ref_nodes", C.c_uint64), ("n_fast_nodes", C.c_uint64), ("n_mats", C.c_uint32),
                ("n_lights", C.c_uint32), ("ref_depth", C.c_uint32), ("fast_depth", C.c_uint32), ("source_key", C.c_uint64),
                ("payload_hash", C.c_uint64), ("device_bytes", C.c_uint64)]


IMAGE_VERIFY = 1
IMAGE_TRUSTED = 2
RENDER_STATS = 1
RENDER_SERIAL = 2
UPLOAD_FAST_TREE_GPU = 1
UPLOAD_FAST_TREE_HOST = 2
KERNEL_CLASSES = ("generate", "trace", "shade", "connect", "shadow", "trace_exact", "shadow_exact")


class BdptError(RuntimeError):
    pass


def build_library(force=False):
    """Compile libbdpt.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    if force:
        subprocess.check_call(["make", "-C", os.path.join(HERE, "csrc"), "-s", "clean"])
    subprocess.check_call(["make", "-C", os.path.join(HERE, "csrc"), "-s", "-j4"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise BdptError("libbdpt.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                            "there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.bdpt_last_error.restype = C.c_char_p
        L.bdpt_last_error.argtypes = [C.c_void_p]
        L.bdpt_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
        L.bdpt_ctx_destroy.argtypes = [C.c_void_p]
        L.bdpt_ctx_destroy.restype = None
        L.bdpt_scene_upload.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32]
        L.bdpt_scene_upload_ex.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_uint32]
        L.bdpt_scene_set_textures.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        L.bdpt_scene_set_media.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]
        L.bdpt_scene_set_spheres.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.bdpt_bvh_info.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]
        L.bdpt_bvh_leaf_order.argtypes = [C.c_void_p, C.c_void_p]
        L.bdpt_host_bvh_build.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.bdpt_trace.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
        L.bdpt_visible.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int]
        L.bdpt_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.bdpt_film_clear.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.bdpt_film_device_ptr.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        L.bdpt_film_download.argtypes = [C.c_void_p, C.c_void_p]
        L.bdpt_film_reduce.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int]
        L.bdpt_comm_unique_id.argtypes = [C.c_void_p]
        L.bdpt_comm_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.bdpt_film_reduce_comm.argtypes = [C.c_void_p, C.c_int]
        L.bdpt_comm_attach.argtypes = [C.c_void_p]
        L.bdpt_render_progress.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.bdpt_bench_read.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.POINTER(C.c_double)]
        L.bdpt_film_size.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.bdpt_test_filter32.argtypes = [C.c_void_p, C.c_double, C.c_void_p, C.c_double, C.c_int, C.POINTER(C.c_float)]
        L.bdpt_film_to_rgb8.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.bdpt_film_to_rgb8_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.bdpt_stats.argtypes = [C.c_void_p, C.c_void_p]
        L.bdpt_stats_reset.argtypes = [C.c_void_p]
        L.bdpt_last_device_ms.argtypes = [C.c_void_p, C.POINTER(C.c_double)]
        L.bdpt_sync.argtypes = [C.c_void_p]
        L.bdpt_kernel_times.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.bdpt_scene_file_load.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        L.bdpt_scene_file_free.argtypes = [C.c_void_p]
        L.bdpt_scene_file_free.restype = None
        L.bdpt_scene_file_info.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64),
                                           C.c_void_p, C.c_char_p, C.c_uint64]
        L.bdpt_scene_file_textures.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.c_void_p, C.POINTER(C.c_uint64), C.c_void_p]
        L.bdpt_scene_file_media.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.c_void_p, C.POINTER(C.c_uint64), C.c_void_p]
        L.bdpt_scene_file_spheres.argtypes = [C.c_void_p, C.POINTER(C.c_uint32), C.c_void_p]
        L.bdpt_scene_file_copy.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.bdpt_write_png.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_int]
        L.bdpt_host_last_error.restype = C.c_char_p
        L.bdpt_scene_image_build.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint32, C.c_uint64,
                                             C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        L.bdpt_scene_image_free.argtypes = [C.c_void_p]
        L.bdpt_scene_image_free.restype = None
        L.bdpt_scene_image_info.argtypes = [C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
        L.bdpt_scene_image_last_error.restype = C.c_char_p
        L.bdpt_scene_upload_image.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32]
        L.bdpt_scene_image_save.argtypes = [C.c_char_p, C.c_void_p, C.c_uint64]
        L.bdpt_scene_image_map.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        L.bdpt_scene_image_unmap.argtypes = [C.c_void_p, C.c_uint64]
        L.bdpt_scene_image_unmap.restype = None
        L.bdpt_scene_source_key.argtypes = [C.c_char_p, C.POINTER(C.c_uint64)]
        L.bdpt_scene_file_open_cached.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                                  C.POINTER(C.c_uint64), C.POINTER(C.c_int)]
        _lib = L
    return _lib


def make_materials(mats):
    """mats: list of (type, (r,g,b), param[, tex]); tex = 1-based index into the texture table, 0 / absent = none"""
    arr = (Material * len(mats))()
    for i, mt in enumerate(mats):
        arr[i].type = int(mt[0])
        arr[i].rgb[:] = [float(x) for x in mt[1]]
        arr[i].param = float(mt[2])
        arr[i].tex = int(mt[3]) if len(mt) > 3 else 0
    return arr


def make_textures(textures):
    """textures: list of dicts {type, a, b, scale, width, height, data} -> (Texture array, uint8 blob); data = the
    noise tables (256 x 3 doubles + 3 x 256 int32, perlin.h:62-66) or the image's width*height*3 bytes."""
    arr = (Texture * max(1, len(textures)))()
    blob = bytearray()
    for i, t in enumerate(textures):
        arr[i].type = int(t["type"])
        arr[i].a[:] = [float(x) for x in t.get("a", (0, 0, 0))]
        arr[i].b[:] = [float(x) for x in t.get("b", (0, 0, 0))]
        arr[i].scale = float(t.get("scale", 1.0))
        arr[i].width, arr[i].height = int(t.get("width", 0)), int(t.get("height", 0))
        while len(blob) % 8:
            blob.append(0)
        arr[i].data_offset = len(blob)
        d = t.get("data")
        if d is not None:
            blob += bytes(d) if isinstance(d, (bytes, bytearray)) else np.ascontiguousarray(d).tobytes()
    return arr, np.frombuffer(bytes(blob) + bytes(8), dtype=np.uint8).copy()


def image_height(width, aspect_ratio):
    """The reference's rule (camera.h:161-162), which bdpt_render applies: the film has this many rows."""
    return max(1, int(width / aspect_ratio))


def aspect_for(width, height):
    """width / height (scene_loader.h:445) -- nudged down by ulps for the ~2 % of (W, H) pairs, e.g. 7x100 or
    23x480, where int(W / (W / H)) truncates to H - 1, so that the reference's rule yields `height` rows."""
    a = float(width) / float(height)
    while image_height(width, a) != int(height):
        a = float(np.nextafter(a, 0.0))
    return a


def make_camera(width, height, spp, depth, lookfrom, lookat, vfov=90.0, vup=(0, 1, 0), background=(0, 0, 0),
                focus_dist=10.0):
    c = Camera()
    c.aspect_ratio = aspect_for(width, height)
    c.image_width, c.samples_per_pixel, c.max_depth = int(width), int(spp), int(depth)
    c.background[:] = background
    c.vfov = vfov
    c.lookfrom[:] = lookfrom
    c.lookat[:] = lookat
    c.vup[:] = vup
    c.defocus_angle = 0.0
    c.focus_dist = focus_dist
    return c


def pack_rays(rays8):
    """(n,8) float64 [ox,oy,oz,dx,dy,dz,tmin,tmax] -> bdpt_ray records (same memory layout)."""
    r = np.ascontiguousarray(rays8, dtype=np.float64).reshape(-1, 8)
    return r.view(RAY_DTYPE).reshape(-1)


def host_bvh_build(verts, fast=False):
    """Reference-topology BVH on the host only: returns (leaf_order, n_ref_nodes, n_fast_nodes|None)."""
    v = np.ascontiguousarray(verts, dtype=np.float64).reshape(-1, 9)
    order = np.empty(len(v), np.uint32)
    nr, nf = C.c_uint64(0), C.c_uint64(0)
    rc = lib().bdpt_host_bvh_build(v.ctypes.data, len(v), order.ctypes.data, C.byref(nr), C.byref(nf) if fast else None)
    if rc:
        raise BdptError("bdpt_host_bvh_build failed: %d %s" % (rc, lib().bdpt_last_error(None).decode()))
    return order, nr.value, (nf.value if fast else None)


class SceneImage:
    """A scene image (include/bdpt.h): flattened soup + both BVHs as one block in the kernels' layout.
    Built on the host (`SceneImage.build`), read from a file (`SceneImage.map`) or wrapped around caller bytes
    (`SceneImage.from_bytes`, e.g. a pinned buffer); uploaded with Context.upload_image."""

    def __init__(self, ptr, nbytes, kind, keep=None):
        self.ptr, self.nbytes, self._kind, self._keep = ptr, int(nbytes), kind, keep

    @classmethod
    def build(cls, verts, mat_id, mats, source_key=0):
        v = np.ascontiguousarray(verts, dtype=np.float64).reshape(-1, 9)
        m = np.ascontiguousarray(mat_id, dtype=np.uint32)
        if len(m) != len(v):
            raise ValueError("mat_id length")
        marr = make_materials(mats) if isinstance(mats, list) else mats
        p, n = C.c_void_p(), C.c_uint64(0)
        rc = lib().bdpt_scene_image_build(v.ctypes.data if len(v) else None, m.ctypes.data if len(v) else None, len(v), C.byref(marr),
                                          len(marr), int(source_key), C.byref(p), C.byref(n))
        if rc:
            raise BdptError("bdpt_scene_image_build failed (%d): %s" % (rc, lib().bdpt_scene_image_last_error().decode()))
        return cls(p.value, n.value, "built")

    @classmethod
    def map(cls, path):
        p, n = C.c_void_p(), C.c_uint64(0)
        if lib().bdpt_scene_image_map(os.fsencode(path), C.byref(p), C.byref(n)):
            raise BdptError(lib().bdpt_host_last_error().decode())
        return cls(p.value, n.value, "mapped")

    @classmethod
    def from_bytes(cls, arr):
        a = np.ascontiguousarray(arr).view(np.uint8).reshape(-1)
        return cls(a.ctypes.data, a.nbytes, "borrowed", keep=a)

    def info(self, verify=False, validate=False):
        i = SceneImageInfo()
        if lib().bdpt_scene_image_info(C.c_void_p(self.ptr), self.nbytes, (1 if verify else 0) | (2 if validate else 0), C.byref(i)):
            raise BdptError(lib().bdpt_scene_image_last_error().decode())
        return {n: int(getattr(i, n)) for n, _ in SceneImageInfo._fields_}

    def bytes(self):
        """A numpy copy of the image bytes."""
        return np.ctypeslib.as_array(C.cast(C.c_void_p(self.ptr), C.POINTER(C.c_uint8)), shape=(self.nbytes,)).copy()

    def save(self, path):
        if lib().bdpt_scene_image_save(os.fsencode(path), C.c_void_p(self.ptr), self.nbytes):
            raise BdptError(lib().bdpt_host_last_error().decode())

    def close(self):
        if self._kind == "built" and self.ptr:
            lib().bdpt_scene_image_free(C.c_void_p(self.ptr))
        elif self._kind == "mapped" and self.ptr:
            lib().bdpt_scene_image_unmap(C.c_void_p(self.ptr), self.nbytes)
        self.ptr, self._keep = None, None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def scene_source_key(yaml_path):
    k = C.c_uint64(0)
    if lib().bdpt_scene_source_key(os.fsencode(yaml_path), C.byref(k)):
        raise BdptError(lib().bdpt_host_last_error().decode())
    return k.value


class CachedSceneFile:
    """bdpt_scene_file_open_cached: camera + scene image of a YAML scene, from `cache_path` when its key matches the
    sources (`hit`), else loaded, built and written there."""

    def __init__(self, yaml_path, cache_path):
        self._h = C.c_void_p()
        p, n, hit = C.c_void_p(), C.c_uint64(0), C.c_int(0)
        if lib().bdpt_scene_file_open_cached(os.fsencode(yaml_path), os.fsencode(cache_path), C.byref(self._h), C.byref(p), C.byref(n),
                                             C.byref(hit)):
            raise BdptError("Failed to load scene: " + lib().bdpt_host_last_error().decode())
        self.hit = bool(hit.value)
        self.image = SceneImage(p.value, n.value, "borrowed", keep=self)
        nt, nm, nl = C.c_uint64(0), C.c_uint32(0), C.c_uint64(0)
        self.camera = Camera()
        name = C.create_string_buffer(512)
        lib().bdpt_scene_file_info(self._h, C.byref(nt), C.byref(nm), C.byref(nl), C.byref(self.camera), name, 512)
        self.n_tris, self.n_mats, self.n_lights, self.output = nt.value, nm.value, nl.value, name.value.decode()

    def close(self):
        if self._h:
            self.image.ptr = None
            lib().bdpt_scene_file_free(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


class SceneFile:
    """load_scene_from_yaml (scene_loader.h:480-523) through the C-ABI: camera fields + flattened triangle soup."""

    def __init__(self, yaml_path):
        h = C.c_void_p()
        rc = lib().bdpt_scene_file_load(os.fsencode(yaml_path), C.byref(h))
        if rc:
            raise BdptError("Failed to load scene: " + lib().bdpt_host_last_error().decode())
        try:
            nt, nm, nl = C.c_uint64(0), C.c_uint32(0), C.c_uint64(0)
            self.camera = Camera()
            name = C.create_string_buffer(512)
            lib().bdpt_scene_file_info(h, C.byref(nt), C.byref(nm), C.byref(nl), C.byref(self.camera), name, 512)
            self.n_lights = nl.value
            self.output = name.value.decode()
            self.verts = np.empty((nt.value, 9), np.float64)
            self.mat_id = np.empty(nt.value, np.uint32)
            self.materials = (Material * max(1, nm.value))()
            lib().bdpt_scene_file_copy(h, self.verts.ctypes.data, self.mat_id.ctypes.data, C.byref(self.materials))
            self.n_mats = nm.value
            nt_, nb_ = C.c_uint32(0), C.c_uint64(0)
            lib().bdpt_scene_file_textures(h, C.byref(nt_), None, C.byref(nb_), None)
            tarr = (Texture * max(1, nt_.value))()
            blob = np.zeros(max(8, nb_.value), np.uint8)
            lib().bdpt_scene_file_textures(h, None, C.byref(tarr), None, blob.ctypes.data)
            self.textures = []   # dicts as Context.set_textures / make_textures take them
            for k in range(nt_.value):
                t = tarr[k]
                n = (256 * 3 * 8 + 3 * 256 * 4) if t.type == TEX_NOISE else (t.width * t.height * 3 if t.type == TEX_IMAGE else 0)
                self.textures.append(dict(type=int(t.type), a=tuple(t.a), b=tuple(t.b), scale=float(t.scale), width=int(t.width),
                                          height=int(t.height), data=blob[t.data_offset:t.data_offset + n].copy() if n else None))
            ns_ = C.c_uint32(0)
            lib().bdpt_scene_file_spheres(h, C.byref(ns_), None)
            sarr = (Sphere * max(1, ns_.value))()
            lib().bdpt_scene_file_spheres(h, None, C.byref(sarr))
            self.spheres = [(tuple(sarr[k].center), float(sarr[k].radius), int(sarr[k].mat_id)) for k in range(ns_.value)]   # as Context.set_spheres takes them
            nm_, nbt = C.c_uint32(0), C.c_uint64(0)
            if lib().bdpt_scene_file_media(h, C.byref(nm_), None, C.byref(nbt), None):
                raise BdptError("Failed to load scene: " + lib().bdpt_host_last_error().decode())
            marr = (Medium * max(1, nm_.value))()
            bverts = np.zeros((max(1, nbt.value), 9), np.float64)
            lib().bdpt_scene_file_media(h, None, C.byref(marr), None, bverts.ctypes.data)
            self.media = [(float(marr[k].density), tuple(marr[k].albedo), bverts[marr[k].first_tri:marr[k].first_tri + marr[k].n_tris].copy())
                          for k in range(nm_.value)]   # as Context.set_media takes them
        finally:
            lib().bdpt_scene_file_free(h)

    def material_list(self):
        """(type, rgb, param) per material; a textured one carries a 4th entry, its 1-based texture index."""
        return [(int(m.type), tuple(m.rgb), float(m.param)) + ((int(m.tex),) if m.tex else ())
                for m in list(self.materials)[:self.n_mats]]

    def camera_with(self, width=None, height=None, spp=None, depth=None):
        """The CLI's --res/--spp/--depth overrides (BASELINE.json configs differ from the YAML values)."""
        c = Camera.from_buffer_copy(self.camera)
        if width and height:
            c.image_width = int(width)
            c.aspect_ratio = aspect_for(width, height)
        if spp:
            c.samples_per_pixel = int(spp)
        if depth is not None:
            c.max_depth = int(depth)
        return c


def write_png(path, rgb):
    rgb = np.ascontiguousarray(rgb, dtype=np.uint8)
    h, w, _ = rgb.shape
    if lib().bdpt_write_png(os.fsencode(path), rgb.ctypes.data, w, h):
        raise BdptError(lib().bdpt_host_last_error().decode())


class Context:
    """One CUDA device. Mirrors the reference's call sequence (main.cpp:62-78):
    upload scene (= load + bvh_node ctor), then trace()/visible() (= world.hit) or render() (= camera::render)."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        rc = lib().bdpt_ctx_create(int(device), C.byref(self._h))
        if rc:
            raise BdptError("bdpt_ctx_create failed (%d): %s — no CPU fallback exists"
                            % (rc, lib().bdpt_last_error(None).decode()))
        self.n_tris = 0
        self.film_shape = None

    def _chk(self, rc, what):
        if rc:
            raise BdptError("%s failed (%d): %s" % (what, rc, lib().bdpt_last_error(self._h).decode()))

    def close(self):
        if self._h:
            lib().bdpt_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def upload(self, verts, mat_id, mats, gpu_tree=None):
        """bdpt_scene_upload; gpu_tree: where the BVH_FAST tree is built -- None: the library's default (the device), True:
        the device (UPLOAD_FAST_TREE_GPU), False: the host's binned-SAH build (UPLOAD_FAST_TREE_HOST)."""
        v = np.ascontiguousarray(verts, dtype=np.float64).reshape(-1, 9)
        m = np.ascontiguousarray(mat_id, dtype=np.uint32)
        if len(m) != len(v):
            raise ValueError("mat_id length")
        marr = make_materials(mats) if isinstance(mats, list) else mats
        self._chk(lib().bdpt_scene_upload_ex(self._h, v.ctypes.data, m.ctypes.data, len(v), C.byref(marr), len(marr),
                                             0 if gpu_tree is None else (UPLOAD_FAST_TREE_GPU if gpu_tree else UPLOAD_FAST_TREE_HOST)), "bdpt_scene_upload")
        self.n_tris = len(v)
        return self

    def set_textures(self, textures):
        """bdpt_scene_set_textures: the table bdpt_material.tex refers to (list of dicts as make_textures takes; [] drops it)."""
        arr, blob = make_textures(textures)
        self._chk(lib().bdpt_scene_set_textures(self._h, C.byref(arr) if textures else None, len(textures),
                                                blob.ctypes.data if textures else None, len(blob) if textures else 0), "bdpt_scene_set_textures")
        return self

    def upload_image(self, image, verify=None, trusted=None):
        """bdpt_scene_upload_image: one allocation + one host-to-device copy of a prebuilt SceneImage.  verify: check the
        payload hash first (default: for images mapped from files); trusted: skip the range check of the payload's
        links and indices (default: only for images built by this process)."""
        kind = getattr(image, "_kind", "borrowed")
        verify = (kind == "mapped") if verify is None else verify
        trusted = (kind == "built") if trusted is None else trusted
        self._chk(lib().bdpt_scene_upload_image(self._h, C.c_void_p(image.ptr), image.nbytes,
                                                (IMAGE_VERIFY if verify else 0) | (IMAGE_TRUSTED if trusted else 0)),
                  "bdpt_scene_upload_image")
        self.n_tris = image.info()["n_tris"]
        return self

    def bvh_info(self, bvh=BVH_REFERENCE):
        n, d = C.c_uint64(0), C.c_uint32(0)
        self._chk(lib().bdpt_bvh_info(self._h, bvh, C.byref(n), C.byref(d)), "bdpt_bvh_info")
        return n.value, d.value

    def leaf_order(self):
        out = np.empty(self.n_tris, np.uint32)
        self._chk(lib().bdpt_bvh_leaf_order(self._h, out.ctypes.data), "bdpt_bvh_leaf_order")
        return out

    def trace(self, rays8, bvh=BVH_REFERENCE, any_hit=False, stats=False, split=False, out=None):
        """Host buffers in, host buffers out (the e2e path). Returns structured HIT_DTYPE array.  `out`: a HIT_DTYPE
        array of len(rays) to fill (e.g. a view of pinned memory; with pinned rays too the library's chunked copies
        overlap the kernels)."""
        r = pack_rays(rays8)
        if out is None:
            out = np.empty(len(r), HIT_DTYPE)
        elif out.dtype != HIT_DTYPE or len(out) != len(r) or not out.flags.c_contiguous:
            raise ValueError("trace(out=...): need a C-contiguous HIT_DTYPE array of len(rays)")
        flags = (TRACE_ANY if any_hit else 0) | (TRACE_STATS if stats else 0) | (TRACE_SPLIT if split else 0)
        self._chk(lib().bdpt_trace(self._h, bvh, r.ctypes.data if len(r) else None, len(r),
                                   out.ctypes.data if len(r) else None, flags), "bdpt_trace")
        return out

    def trace_device(self, d_rays, n, d_out, bvh=BVH_REFERENCE, any_hit=False, stats=False):
        """Device pointers (ints) in and out; nothing is copied."""
        flags = TRACE_DEVICE_PTRS | (TRACE_ANY if any_hit else 0) | (TRACE_STATS if stats else 0)
        self._chk(lib().bdpt_trace(self._h, bvh, C.c_void_p(d_rays), n, C.c_void_p(d_out), flags), "bdpt_trace")

    def visible(self, a, b, bvh=BVH_REFERENCE):
        a = np.ascontiguousarray(a, dtype=np.float64).reshape(-1, 3)
        b = np.ascontiguousarray(b, dtype=np.float64).reshape(-1, 3)
        out = np.empty(len(a), np.uint8)
        self._chk(lib().bdpt_visible(self._h, bvh, a.ctypes.data, b.ctypes.data, len(a), out.ctypes.data, 0), "bdpt_visible")
        return out.astype(bool)

    def set_spheres(self, spheres):
        """bdpt_scene_set_spheres: analytic sphere prims (an extension; the reference tessellates).  spheres: list of
        ((cx, cy, cz), radius, material index into the uploaded scene's table); [] drops them."""
        arr = (Sphere * max(1, len(spheres)))()
        for k, (c, r, m) in enumerate(spheres):
            arr[k].center[:] = [float(x) for x in c]
            arr[k].radius, arr[k].mat_id = float(r), int(m)
        self._chk(lib().bdpt_scene_set_spheres(self._h, C.byref(arr) if spheres else None, len(spheres)), "bdpt_scene_set_spheres")

    def set_media(self, media):
        """bdpt_scene_set_media: constant_medium + isotropic (constant_medium.h:8-61, material.h:149-171).  media: list of
        (density, (r, g, b), boundary triangles n x 9), closed convex meshes; [] drops them."""
        arr = (Medium * max(1, len(media)))()
        soup, first = [], 0
        for k, (density, rgb, tris) in enumerate(media):
            t = np.ascontiguousarray(tris, dtype=np.float64).reshape(-1, 9)
            arr[k].density = float(density)
            arr[k].albedo[:] = [float(x) for x in rgb]
            arr[k].first_tri, arr[k].n_tris = first, len(t)
            soup.append(t); first += len(t)
        b = np.ascontiguousarray(np.concatenate(soup) if soup else np.zeros((0, 9)), dtype=np.float64)
        self._chk(lib().bdpt_scene_set_media(self._h, C.byref(arr) if media else None, len(media), b.ctypes.data if len(b) else None, len(b)),
                  "bdpt_scene_set_media")

    def film_clear(self, width, height):
        self._chk(lib().bdpt_film_clear(self._h, int(width), int(height)), "bdpt_film_clear")
        self.film_shape = (int(height), int(width), 3)

    def render(self, cam, seed=0, bvh=BVH_REFERENCE, strata=(0, 0, 1), rows=(0, 0), batch_samples=0, clear=True,
               height=None, stats=False, streams=0, serial=False, weighting=WEIGHT_REF_SUM):
        """camera::render's sample loop; returns nothing — read the film with film().  The film's row count is
        the reference's int(image_width / aspect_ratio) (camera.h:161-162), decided by the library; `height`, when
        given, is only checked against it (make_camera / camera_with pick an aspect ratio that yields it)."""
        W = cam.image_width
        H = image_height(W, cam.aspect_ratio)
        if height is not None and int(height) != H:
            raise ValueError("height=%d, but image_width %d / aspect_ratio %r gives %d rows (camera.h:161); build the camera "
                             "with make_camera(width, height, ...)" % (height, W, cam.aspect_ratio, H))
        if clear or self.film_shape != (H, W, 3):
            self.film_clear(W, H)
        o = RenderOpts()
        o.seed = seed
        o.stratum_begin, o.stratum_end, o.stratum_step = strata
        o.bvh = bvh
        o.row_begin, o.row_end = rows
        o.batch_samples = batch_samples
        o.flags = (RENDER_STATS if stats else 0) | (RENDER_SERIAL if serial else 0)
        o.streams = streams
        o.weighting = weighting
        self._chk(lib().bdpt_render(self._h, C.byref(cam), C.byref(o)), "bdpt_render")
        if self.film_size() != (W, H):   # one source of truth: the library's film
            raise BdptError("film is %r after bdpt_render, expected %r" % (self.film_size(), (W, H)))

    def film_size(self):
        w, h = C.c_int(0), C.c_int(0)
        self._chk(lib().bdpt_film_size(self._h, C.byref(w), C.byref(h)), "bdpt_film_size")
        return w.value, h.value

    def film(self, out=None):
        """The fp64 RGB sum film (camera.h:55) copied to the host; `out`: a C-contiguous float64 array of
        film_shape to fill (e.g. a view of pinned memory, reused from frame to frame)."""
        if out is None:
            out = np.empty(self.film_shape, np.float64)
        elif out.dtype != np.float64 or tuple(out.shape) != tuple(self.film_shape) or not out.flags.c_contiguous:
            raise ValueError("film(out=...): need a C-contiguous float64 array of shape %r" % (tuple(self.film_shape),))
        self._chk(lib().bdpt_film_download(self._h, out.ctypes.data), "bdpt_film_download")
        return out

    def film_device_ptr(self):
        p, n = C.c_void_p(), C.c_uint64(0)
        self._chk(lib().bdpt_film_device_ptr(self._h, C.byref(p), C.byref(n)), "bdpt_film_device_ptr")
        return p.value, n.value

    def film_rgb8(self, spp):
        out = np.empty(self.film_shape, np.uint8)
        self._chk(lib().bdpt_film_to_rgb8(self._h, int(spp), out.ctypes.data), "bdpt_film_to_rgb8")
        return out

    def film_rgb8_noalloc(self, spp, out=None):
        """bdpt_film_to_rgb8_noalloc: the same bytes through the library's static device scratch (no cudaMalloc);
        films beyond NOALLOC_RGB8_MAX_BYTES are refused."""
        if out is None:
            out = np.empty(self.film_shape, np.uint8)
        self._chk(lib().bdpt_film_to_rgb8_noalloc(self._h, int(spp), out.ctypes.data), "bdpt_film_to_rgb8_noalloc")
        return out

    def stats(self):
        s = Stats()
        self._chk(lib().bdpt_stats(self._h, C.byref(s)), "bdpt_stats")
        return {n: int(getattr(s, n)) for n, _ in Stats._fields_}

    def stats_reset(self):
        self._chk(lib().bdpt_stats_reset(self._h), "bdpt_stats_reset")

    def kernel_times(self):
        """{class: (ms, launches)} of bdpt_render's kernels since the last stats_reset (CUDA events on the ctx stream)."""
        ms = (C.c_double * len(KERNEL_CLASSES))()
        n = (C.c_uint64 * len(KERNEL_CLASSES))()
        self._chk(lib().bdpt_kernel_times(self._h, ms, n), "bdpt_kernel_times")
        return {k: (ms[i], int(n[i])) for i, k in enumerate(KERNEL_CLASSES)}

    def last_device_ms(self):
        ms = C.c_double(0)
        self._chk(lib().bdpt_last_device_ms(self._h, C.byref(ms)), "bdpt_last_device_ms")
        return ms.value

    def sync(self):
        self._chk(lib().bdpt_sync(self._h), "bdpt_sync")

    def bench_read(self, nbytes, reps=20):
        """Read bandwidth in GB/s of an nbytes device buffer (bdpt_bench_read): L2-resident when it fits the L2."""
        g = C.c_double(0)
        self._chk(lib().bdpt_bench_read(self._h, int(nbytes), int(reps), C.byref(g)), "bdpt_bench_read")
        return g.value

    def comm_init(self, nranks, rank, unique_id):
        """Join the NCCL communicator of the film reduce (one process per GPU); unique_id: the 128 bytes of
        comm_unique_id() from one rank, carried to the others by the launcher (torch.distributed, MPI, a file)."""
        buf = (C.c_char * 128).from_buffer_copy(bytes(unique_id))
        self._chk(lib().bdpt_comm_init(self._h, int(nranks), int(rank), buf), "bdpt_comm_init")

    def comm_attach(self):
        """Take over the communicator an earlier context of this process joined on the same device."""
        self._chk(lib().bdpt_comm_attach(self._h), "bdpt_comm_attach")

    def film_reduce(self, root=0):
        """The one exchange step of a frame: NCCL sum-reduce of the ranks' device films onto `root`, in place."""
        self._chk(lib().bdpt_film_reduce_comm(self._h, int(root)), "bdpt_film_reduce_comm")
        return self.last_device_ms()

    def progress(self):
        d, t = C.c_uint64(0), C.c_uint64(0)
        self._chk(lib().bdpt_render_progress(self._h, C.byref(d), C.byref(t)), "bdpt_render_progress")
        return d.value, t.value


# ---------------------------------------------------------------- multi-GPU: stratum sharding
def comm_unique_id():
    """128 bytes naming a new NCCL communicator (bdpt_comm_unique_id): create on one rank, send to the others."""
    buf = (C.c_char * 128)()
    if lib().bdpt_comm_unique_id(buf):
        raise BdptError("bdpt_comm_unique_id failed (is libnccl.so.2 loadable?)")
    return bytes(buf)


def film_reduce(contexts, root=0):
    """bdpt_film_reduce: one process, several devices -- sums the contexts' films onto contexts[root] over NCCL."""
    arr = (C.c_void_p * len(contexts))(*[c._h for c in contexts])
    rc = lib().bdpt_film_reduce(arr, len(contexts), int(root))
    if rc:
        raise BdptError("bdpt_film_reduce failed (%d): %s" % (rc, lib().bdpt_last_error(contexts[root]._h).decode()))
    return contexts[root].last_device_ms()


def shard_strata(rank, world, samples_per_pixel):
    """The strata rank `rank` of `world` renders: (begin, end, step) for bdpt_render_opts — stratum s belongs to
    rank s mod world, which keeps the reference's stratification (camera.h:164,213-223) and gives every rank the
    same share of every pixel."""
    n = int(np.sqrt(float(samples_per_pixel)))
    if not 0 <= rank < world:
        raise ValueError("rank %d outside world %d" % (rank, world))
    return (int(rank), n * n, int(world))


class CudaArray:
    """__cuda_array_interface__ view of a device pointer owned by libbdpt (the film), so that torch.distributed
    can reduce it in place over NCCL."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 3}


def reduce_film(film, dist, dst=0):
    """The one exchange step of a frame: SUM of the ranks' film sums onto `dst` (NCCL on device tensors, gloo on
    CPU tensors). `film` is a torch tensor and is reduced in place."""
    dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return film


def join_comm(ctx, dist):
    """bdpt_comm_init for the ranks of an initialised torch.distributed group: rank 0 creates the id, the group's
    own broadcast carries it (plumbing only; the reduce itself is the library's)."""
    rank, world = dist.get_rank(), dist.get_world_size()
    box = [comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    ctx.comm_init(world, rank, box[0])


def render_sharded(ctx, cam, dist=None, seed=0, bvh=BVH_FAST, height=None, device=None):
    """camera::render across the ranks of an initialised torch.distributed group: every rank renders its strata
    into its own film, then ONE film reduce inside the library (bdpt_film_reduce_comm; the context must have joined
    the communicator, join_comm). Returns the summed film (numpy, H x W x 3) on rank 0, None elsewhere.
    With dist=None it is a plain single-GPU render."""
    rank, world = (dist.get_rank(), dist.get_world_size()) if dist is not None else (0, 1)
    ctx.render(cam, seed=seed, bvh=bvh, strata=shard_strata(rank, world, cam.samples_per_pixel), clear=True, height=height)
    if dist is not None and world > 1:
        ctx.film_reduce(root=0)
    return ctx.film() if rank == 0 else None
