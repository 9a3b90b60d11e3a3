"""Small procedural triangle scenes for parity tests (numpy only; no oracle, no product imports).

Every scene is (verts[N,9] float64, mat_id[N] uint32, mats list of (type,(r,g,b),param), camera kwargs).
Material types follow include/bdpt.h / oracle_api.h: 0 lambertian, 1 metal, 2 dielectric, 3 light.
"""
import numpy as np

LAMB, METAL, GLASS, LIGHT = 0, 1, 2, 3


def quad(q, u, v):
    """Two triangles in the reference's add_quad_triangles order (triangle.h:232-241)."""
    q, u, v = (np.asarray(x, dtype=np.float64) for x in (q, u, v))
    return np.array([np.concatenate([q, q + u, q + v]), np.concatenate([q + u, q + u + v, q + v])])


def box(a, b, rot_y_deg=0.0, translate=(0, 0, 0)):
    a, b = np.minimum(a, b).astype(np.float64), np.maximum(a, b).astype(np.float64)
    c = lambda i, j, k: np.array([(a, b)[i][0], (a, b)[j][1], (a, b)[k][2]])
    faces = [(c(0, 0, 1), c(1, 0, 1), c(1, 1, 1)), (c(0, 0, 1), c(1, 1, 1), c(0, 1, 1)),
             (c(0, 0, 0), c(0, 1, 0), c(1, 1, 0)), (c(0, 0, 0), c(1, 1, 0), c(1, 0, 0)),
             (c(0, 0, 0), c(0, 0, 1), c(0, 1, 1)), (c(0, 0, 0), c(0, 1, 1), c(0, 1, 0)),
             (c(1, 0, 1), c(1, 0, 0), c(1, 1, 0)), (c(1, 0, 1), c(1, 1, 0), c(1, 1, 1)),
             (c(0, 1, 1), c(1, 1, 1), c(1, 1, 0)), (c(0, 1, 1), c(1, 1, 0), c(0, 1, 0)),
             (c(0, 0, 0), c(1, 0, 0), c(1, 0, 1)), (c(0, 0, 0), c(1, 0, 1), c(0, 0, 1))]
    th = np.deg2rad(rot_y_deg)
    s, co = np.sin(th), np.cos(th)
    out = []
    for f in faces:
        tri = []
        for p in f:
            if rot_y_deg != 0.0:
                p = np.array([co * p[0] + s * p[2], p[1], -s * p[0] + co * p[2]])
            tri.append(p + np.asarray(translate, dtype=np.float64))
        out.append(np.concatenate(tri))
    return np.array(out)


def uv_sphere(center, radius, lat_steps=16, lon_steps=32):
    """Same tessellation rule as the reference's `type: Sphere` (scene_loader.h:212-242)."""
    center = np.asarray(center, dtype=np.float64)

    def pt(theta, phi):
        st = np.sin(theta)
        return center + np.array([radius * st * np.cos(phi), radius * np.cos(theta), radius * st * np.sin(phi)])

    tris = []
    for lat in range(lat_steps):
        t0, t1 = np.pi * (lat / lat_steps), np.pi * ((lat + 1) / lat_steps)
        for lon in range(lon_steps):
            p0, p1 = 2.0 * np.pi * (lon / lon_steps), 2.0 * np.pi * ((lon + 1) / lon_steps)
            p00, p01, p10, p11 = pt(t0, p0), pt(t0, p1), pt(t1, p0), pt(t1, p1)
            if lat > 0:
                tris.append(np.concatenate([p00, p10, p11]))
            if lat < lat_steps - 1:
                tris.append(np.concatenate([p00, p11, p01]))
    return np.array(tris)


class SceneBuilder:
    def __init__(self):
        self.verts, self.mat_id, self.mats = [], [], []

    def material(self, t, rgb=(0, 0, 0), param=0.0):
        self.mats.append((t, tuple(float(x) for x in rgb), float(param)))
        return len(self.mats) - 1

    def add(self, tris, m):
        tris = np.asarray(tris, dtype=np.float64).reshape(-1, 9)
        self.verts.append(tris)
        self.mat_id.append(np.full(len(tris), m, dtype=np.uint32))

    def done(self):
        return np.concatenate(self.verts), np.concatenate(self.mat_id), self.mats


def cornell(spheres=True, sphere_steps=(16, 32), f32_round=False):
    """Closed Cornell-like room (555 units) with a light quad, a rotated lambertian box and, optionally,
    a glass and a fuzzy-metal UV sphere.  24 + 12 (+ 2*960) triangles."""
    sb = SceneBuilder()
    red = sb.material(LAMB, (.65, .05, .05)); white = sb.material(LAMB, (.73, .73, .73))
    green = sb.material(LAMB, (.12, .45, .15)); light = sb.material(LIGHT, (15, 15, 15))
    sb.add(quad((555, 0, 0), (0, 0, 555), (0, 555, 0)), green)
    sb.add(quad((0, 0, 555), (0, 0, -555), (0, 555, 0)), red)
    sb.add(quad((0, 555, 0), (555, 0, 0), (0, 0, 555)), white)
    sb.add(quad((0, 0, 555), (555, 0, 0), (0, 0, -555)), white)
    sb.add(quad((555, 0, 555), (-555, 0, 0), (0, 555, 0)), white)
    sb.add(quad((213, 554, 227), (130, 0, 0), (0, 0, 105)), light)  # normal points down (-y)
    sb.add(box((0, 0, 0), (165, 330, 165), 15.0, (265, 0, 295)), white)
    if spheres:
        glass = sb.material(GLASS, (1, 1, 1), 1.5); metal = sb.material(METAL, (.8, .85, .88), 0.15)
        sb.add(uv_sphere((190, 90, 190), 90, *sphere_steps), glass)
        sb.add(uv_sphere((400, 70, 120), 70, *sphere_steps), metal)
    v, m, mats = sb.done()
    if f32_round:
        v = v.astype(np.float32).astype(np.float64)
    cam = dict(lookfrom=(278, 278, -800), lookat=(278, 278, 0), vfov=40.0, background=(0, 0, 0))
    return v, m, mats, cam


TEX_SOLID, TEX_CHECKER, TEX_NOISE, TEX_IMAGE = 0, 1, 2, 3


def cornell_textured(noise=True):
    """The Cornell room with the reference's texture classes on it (texture.h:28-87): a checker floor, an
    image on the back wall (a 12 x 9 procedural RGB pattern, raw bytes), a checker-textured LIGHT, a Perlin-marble box
    (noise=True) and plain red / green side walls.  Returns (verts, mat_id, mats, cam, textures): mats carry a 4th entry,
    the 1-based texture index; textures are dicts as oracle.make_textures / Context.set_textures take them."""
    sb = SceneBuilder()
    yy, xx = np.mgrid[0:9, 0:12]
    img = np.stack([(xx * 23 + yy * 7) % 256, (xx * xx * 5 + 40 * yy) % 256, (255 - 20 * xx + 3 * yy * yy) % 256], -1).astype(np.uint8)
    textures = [dict(type=TEX_CHECKER, a=(.8, .8, .8), b=(.15, .2, .6), scale=60.0),
                dict(type=TEX_IMAGE, width=12, height=9, data=img),
                dict(type=TEX_CHECKER, a=(15, 15, 15), b=(15, 6, 2), scale=25.0)]
    if noise:
        textures.append(dict(type=TEX_NOISE, scale=0.04))

    def mat(t, rgb, param=0.0, tex=0):
        sb.mats.append((t, tuple(float(x) for x in rgb), float(param), int(tex)))
        return len(sb.mats) - 1

    red = mat(LAMB, (.65, .05, .05)); green = mat(LAMB, (.12, .45, .15)); white = mat(LAMB, (.73, .73, .73))
    floor = mat(LAMB, (.5, .5, .5), tex=1); back = mat(LAMB, (.5, .5, .5), tex=2); light = mat(LIGHT, (15, 15, 15), tex=3)
    boxm = mat(LAMB, (.73, .73, .73), tex=4 if noise else 0)
    sb.add(quad((555, 0, 0), (0, 0, 555), (0, 555, 0)), green)
    sb.add(quad((0, 0, 555), (0, 0, -555), (0, 555, 0)), red)
    sb.add(quad((0, 555, 0), (555, 0, 0), (0, 0, 555)), white)
    sb.add(quad((0, 0, 555), (555, 0, 0), (0, 0, -555)), floor)
    sb.add(quad((555, 0, 555), (-555, 0, 0), (0, 555, 0)), back)
    sb.add(quad((213, 554, 227), (130, 0, 0), (0, 0, 105)), light)
    sb.add(box((0, 0, 0), (165, 330, 165), 15.0, (265, 0, 295)), boxm)
    v, m, mats = sb.done()
    cam = dict(lookfrom=(278, 278, -800), lookat=(278, 278, 0), vfov=40.0, background=(0, 0, 0))
    return v, m, mats, cam, textures


def unit_room(seed=1, n_blobs=6):
    """1-unit-scale open scene (coffee_machine-like scale): floor, three light quads, random small
    lambertian/metal/glass spheres; non-black background so misses contribute."""
    rng = np.random.default_rng(seed)
    sb = SceneBuilder()
    floor = sb.material(LAMB, (147 / 255,) * 3); light = sb.material(LIGHT, (245, 245, 245))
    sb.add(quad((-1, 0, -1), (0, 0, 2), (2, 0, 0)), floor)
    sb.add(quad((-0.36, 0.45, -0.01), (0.16, 0, 0.35), (0, -0.449, 0)), light)
    sb.add(quad((0.32, 0.027, 0.229), (0, 0.449, 0), (0.005, 0, -0.365)), light)
    sb.add(quad((0.23, 0.504, 0.267), (-0.46, 0, 0), (0, 0, -0.46)), light)
    kinds = [(LAMB, 0.0), (METAL, 0.1), (GLASS, 1.5), (GLASS, 0.763), (METAL, 0.0), (LAMB, 0.0)]
    for k in range(n_blobs):
        t, p = kinds[k % len(kinds)]
        m = sb.material(t, tuple(rng.uniform(0.2, 0.95, 3)), p)
        c = np.array([rng.uniform(-0.2, 0.2), rng.uniform(0.04, 0.3), rng.uniform(-0.15, 0.2)])
        sb.add(uv_sphere(c, rng.uniform(0.03, 0.07), 8, 12), m)
    v, m, mats = sb.done()
    cam = dict(lookfrom=(-0.003, 0.198, 0.828), lookat=(0.0029, 0.1666, 0.0227), vfov=30.0, background=(0.05, 0.07, 0.1))
    return v, m, mats, cam


def random_soup(n, seed=0, extent=10.0, size=1.0):
    rng = np.random.default_rng(seed)
    c = rng.uniform(-extent, extent, (n, 1, 3))
    v = (c + rng.uniform(-size, size, (n, 3, 3))).reshape(n, 9)
    v = v.astype(np.float32).astype(np.float64)
    return v, np.zeros(n, np.uint32), [(LAMB, (0.5, 0.5, 0.5), 0.0)]


def ray_batch(verts, n, seed=0x5EED0001, cam=None, width=64, height=64):
    """Three equal classes of rays, values fp32-representable:
    (A) camera rays with un-normalised directions, (B) incoherent rays in the inflated scene AABB,
    (C) shadow-style segments between two surface points, t in [0.001, dist-0.001]."""
    rng = np.random.default_rng(seed)
    P = verts.reshape(-1, 3, 3)
    lo, hi = P.min(axis=(0, 1)), P.max(axis=(0, 1))
    ext = hi - lo
    nA = n // 3 if cam is not None else 0
    nC = n // 3
    nB = n - nA - nC
    out = []
    if nA:
        o = np.asarray(cam["lookfrom"], dtype=np.float64)
        at = np.asarray(cam["lookat"], dtype=np.float64)
        w = o - at; w /= np.linalg.norm(w)
        u = np.cross((0, 1, 0), w); u /= np.linalg.norm(u)
        v = np.cross(w, u)
        h = np.tan(np.deg2rad(cam.get("vfov", 90.0)) / 2) * 10.0
        x = rng.uniform(-1, 1, nA) * h * (width / height); y = rng.uniform(-1, 1, nA) * h
        d = x[:, None] * u + y[:, None] * v - 10.0 * w
        r = np.concatenate([np.broadcast_to(o, (nA, 3)), d, np.full((nA, 1), 0.001), np.full((nA, 1), np.inf)], axis=1)
        out.append(r)
    o = rng.uniform(lo - 0.05 * ext, hi + 0.05 * ext, (nB, 3))
    d = rng.normal(size=(nB, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    out.append(np.concatenate([o, d, np.full((nB, 1), 0.001), np.full((nB, 1), np.inf)], axis=1))

    def surf(k):
        area = 0.5 * np.linalg.norm(np.cross(P[:, 1] - P[:, 0], P[:, 2] - P[:, 0]), axis=1)
        idx = rng.choice(len(P), size=k, p=area / area.sum())
        a, b = rng.uniform(size=k), rng.uniform(size=k)
        f = a + b > 1
        a[f], b[f] = 1 - a[f], 1 - b[f]
        return P[idx, 0] + a[:, None] * (P[idx, 1] - P[idx, 0]) + b[:, None] * (P[idx, 2] - P[idx, 0])

    a, b = surf(nC).astype(np.float32).astype(np.float64), surf(nC).astype(np.float32).astype(np.float64)
    d = b - a
    dist = np.linalg.norm(d, axis=1, keepdims=True)
    dist[dist == 0] = 1.0
    d = (d / dist).astype(np.float32).astype(np.float64)
    out.append(np.concatenate([a, d, np.full((nC, 1), 0.001), np.maximum(dist - 0.001, 0.002)], axis=1))
    rays = np.concatenate(out, axis=0)
    rays[:, :6] = rays[:, :6].astype(np.float32).astype(np.float64)
    t = rays[:, 7]
    fin = np.isfinite(t)
    t[fin] = t[fin].astype(np.float32).astype(np.float64)
    return np.ascontiguousarray(rays)


def sphere_grid(g=50, lat_steps=41, lon_steps=50, spacing=2.5, radius=1.0):
    """BASELINE.json config 5: floor + 3 light quads + a g x g grid of UV spheres; sphere i is
    lambertian when i % 10 == 0, else dielectric(1.5) for odd i and metal(0.7, fuzz 0.1) for even i.
    g=50, 41 x 50 steps -> 2,500 spheres x 4,000 triangles = 10,000,008 triangles with floor and lights.
    Camera above the grid looking at its centre."""
    unit = uv_sphere((0, 0, 0), 1.0, lat_steps, lon_steps).reshape(-1, 3, 3)          # T x 3 x 3
    ext = g * spacing
    ij = np.stack(np.meshgrid(np.arange(g), np.arange(g), indexing="ij"), -1).reshape(-1, 2)
    centers = np.stack([(ij[:, 0] + 0.5) * spacing - ext / 2, np.full(len(ij), radius), (ij[:, 1] + 0.5) * spacing - ext / 2], -1)
    sb = SceneBuilder()
    floor = sb.material(LAMB, (.6, .6, .6)); light = sb.material(LIGHT, (40, 40, 40))
    diffuse = sb.material(LAMB, (.7, .3, .2)); glass = sb.material(GLASS, param=1.5); metal = sb.material(METAL, (.7, .7, .7), 0.1)
    m = 0.55 * ext
    sb.add(quad((-m, 0, -m), (0, 0, 2 * m), (2 * m, 0, 0)), floor)                         # normal +y
    h = 0.35 * ext + 4 * radius
    for cx in (-0.3 * ext, 0.0, 0.3 * ext):                                                # three lights, normals -y
        sb.add(quad((cx - 0.08 * ext, h, -0.08 * ext), (0.16 * ext, 0, 0), (0, 0, 0.16 * ext)), light)
    idx = np.arange(len(centers))
    mat_of = np.where(idx % 10 == 0, diffuse, np.where(idx % 2 == 1, glass, metal)).astype(np.uint32)
    verts = (unit[None] * radius + centers[:, None, None, :]).reshape(-1, 9)
    sb.verts.append(verts)
    sb.mat_id.append(np.repeat(mat_of, len(unit)))
    v, mid, mats = sb.done()
    cam = dict(lookfrom=(0.0, 0.55 * ext, -0.75 * ext), lookat=(0.0, 0.0, 0.0), vfov=40.0)
    return v, mid, mats, cam


def box_mesh(lo, hi):
    """The 12 triangles of an axis-aligned box (a closed convex boundary for constant_medium), n x 9."""
    lo = np.asarray(lo, np.float64); hi = np.asarray(hi, np.float64)
    c = np.array([[lo[0] if not i & 1 else hi[0], lo[1] if not i & 2 else hi[1], lo[2] if not i & 4 else hi[2]] for i in range(8)])
    faces = [(0, 1, 3, 2), (4, 6, 7, 5), (0, 4, 5, 1), (2, 3, 7, 6), (0, 2, 6, 4), (1, 5, 7, 3)]
    t = []
    for a, b, cc, d in faces:
        t.append(np.concatenate([c[a], c[b], c[cc]])); t.append(np.concatenate([c[a], c[cc], c[d]]))
    return np.array(t)


def cornell_media():
    """Two overlapping box-shaped media for the Cornell room: a dense white smoke and a thin blue one (density, colour, boundary)."""
    return [(0.01, (0.9, 0.9, 0.9), box_mesh((100, 0, 100), (300, 300, 300))),
            (0.004, (0.2, 0.4, 0.9), box_mesh((250, 150, 50), (500, 400, 250)))]
