"""Analytic sphere prims (bdpt_scene_set_spheres; north_star: "ray-triangle/sphere intersection").  NOT a reference behaviour:
the reference tessellates every sphere into 960 triangles (scene_loader.h:212-242) -- so there is no reference code to pin
this against, and the oracle is the port's own restatement (oracle/port.cpp: sphere_hit), itself checked against the closed
form and against ever finer tessellations of the same sphere traced by the reference-pinned triangle path.
  GPU: film equal to the port's sample for sample (1e-9 of the peak), both tree flavours, alone and together with media."""
import numpy as np
import pytest

from oracle.oracle import Oracle, make_camera as ocam, RNG_PHILOX
from tools import testscenes as ts

LAMB, METAL, GLASS, LIGHT = 0, 1, 2, 3


def rel_err(a, b):
    return float(np.abs(a - b).max() / max(1e-300, np.abs(b).max()))


def room_with_sphere_materials():
    """The Cornell room plus three materials for spheres (glass, fuzzy metal, lambertian); returns the sphere list too."""
    v, m, mats, cam = ts.cornell(False)
    mats = list(mats) + [(GLASS, (1, 1, 1), 1.5), (METAL, (0.8, 0.85, 0.88), 0.1), (LAMB, (0.2, 0.6, 0.3), 0.0)]
    n = len(mats)
    spheres = [((190.0, 90.0, 190.0), 90.0, n - 3), ((400.0, 120.0, 350.0), 120.0, n - 2), ((300.0, 400.0, 300.0), 60.0, n - 1)]
    return v, m, mats, cam, spheres


def test_port_sphere_hit_is_the_closed_form():
    """port_trace asks the sphere side list after the triangles (hit_world's deterministic part): nearer root from outside, farther
    root from inside, misses, the interval's inclusive ends, and a triangle in front wins."""
    c, r = np.array([1.0, 2.0, 3.0]), 2.5
    far = np.array([[100.0, 100, 100, 101, 100, 100, 100, 101, 100]])          # one triangle far away: the scene needs a world
    o = Oracle("port").load(far, np.zeros(1, np.uint32), [(LAMB, (0.5, 0.5, 0.5), 0.0)]).set_spheres([(tuple(c), r, 0)])
    rng = np.random.default_rng(3)
    n = 5000
    org = c + rng.normal(size=(n, 3)) * 6.0
    d = (c + rng.normal(size=(n, 3)) * r * 0.9) - org
    d *= rng.uniform(0.2, 3.0, size=(n, 1))                                      # un-normalised directions
    rays = np.concatenate([org, d, np.full((n, 1), 0.001), np.full((n, 1), np.inf)], axis=1)
    prim, t, _, _ = o.trace(rays)
    a = (d * d).sum(1); h = (d * (c - org)).sum(1); cc = ((c - org) ** 2).sum(1) - r * r
    disc = h * h - a * cc
    sq = np.sqrt(np.maximum(disc, 0))
    t0, t1 = (h - sq) / a, (h + sq) / a
    want = np.where(disc < 0, np.inf, np.where(t0 >= 0.001, t0, np.where(t1 >= 0.001, t1, np.inf)))
    hit = np.isfinite(want)
    assert hit.sum() > 2000 and (~hit).sum() > 10
    assert np.all(prim[hit] == 0x10000000) and np.all(prim[~hit] == 0xFFFFFFFF)
    assert np.abs(t[hit] - want[hit]).max() <= 1e-12 * np.abs(want[hit]).max()
    inside = ((org - c) ** 2).sum(1) < r * r
    assert inside.sum() > 50 and np.all(t[inside & hit] == ((h + sq) / a)[inside & hit])   # bit for bit: the same formula
    # interval ends are inclusive (interval.h:22-24, as triangle.h:68 uses them)
    k = int(np.flatnonzero(hit & ~inside)[0])
    one = rays[k:k + 1].copy(); one[0, 7] = t[k]
    assert o.trace(one)[0][0] == 0x10000000
    one[0, 7] = np.nextafter(t[k], 0)
    assert o.trace(one)[0][0] == 0xFFFFFFFF
    # a triangle in front of the sphere wins
    wall = np.array([[-50.0, -50, -2, 50, -50, -2, 0, 50, -2]])
    o2 = Oracle("port").load(wall, np.zeros(1, np.uint32), [(LAMB, (0.5, 0.5, 0.5), 0.0)]).set_spheres([(tuple(c), r, 0)])
    p2, t2, _, _ = o2.trace(np.array([[1.0, 2.0, -10.0, 0, 0, 1, 0.001, np.inf]]))
    assert p2[0] == 0 and t2[0] == 8.0


def test_port_render_with_analytic_spheres_looks_like_the_tessellated_scene():
    """Same room, same three spheres: analytic prims against 40 x 80-step tessellations (6,240 triangles each).  Means agree
    within Monte-Carlo noise and the faceting error; the sphere-free room is far away."""
    v, m, mats, cam, spheres = room_with_sphere_materials()
    W = H = 24
    spp, depth = 64, 6
    c = ocam(W, H, spp, depth, **cam)
    ana = Oracle("port").load(v, m, mats).set_spheres(spheres).render(c, rng_mode=RNG_PHILOX, seed=5, nthreads=8, height=H)
    tv, tm = [np.asarray(v).reshape(-1, 9)], [np.asarray(m)]
    for ctr, rad, mat in spheres:
        tess = ts.uv_sphere(ctr, rad, 40, 80).reshape(-1, 9)
        tv.append(tess); tm.append(np.full(len(tess), mat, np.uint32))
    tes = Oracle("port").load(np.concatenate(tv), np.concatenate(tm), mats).render(c, rng_mode=RNG_PHILOX, seed=6, nthreads=8, height=H)
    bare = Oracle("port").load(v, m, mats).render(c, rng_mode=RNG_PHILOX, seed=5, nthreads=8, height=H)
    assert abs(ana.mean() / tes.mean() - 1.0) < 0.08
    assert np.mean(np.abs(ana - bare) > 1e-9) > 0.2


@pytest.mark.gpu
@pytest.mark.parametrize("bvh", [0, 1])
def test_sphere_film_equals_the_port_sample_for_sample(bvh, bdpt):
    v, m, mats, cam, spheres = room_with_sphere_materials()
    W, H, spp, depth = 40, 30, 4, 8
    ref = Oracle("port").load(v, m, mats).set_spheres(spheres).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=11, nthreads=8, height=H)
    bare = Oracle("port").load(v, m, mats).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=11, nthreads=8, height=H)
    assert rel_err(bare, ref) > 0.05
    with bdpt.Context(0) as ctx:
        ctx.upload(v, m, mats)
        ctx.set_spheres(spheres)
        ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=11, bvh=bvh, height=H)
        assert rel_err(ctx.film(), ref) < 1e-9
        # together with media
        media = ts.cornell_media()
        both = Oracle("port").load(v, m, mats).set_spheres(spheres).set_media(media).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX,
                                                                                          seed=11, nthreads=8, height=H)
        ctx.set_media(media)
        ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=11, bvh=bvh, height=H)
        assert rel_err(ctx.film(), both) < 1e-9
        ctx.set_media([]); ctx.set_spheres([])
        ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=11, bvh=bvh, height=H)
        assert rel_err(ctx.film(), bare) < 1e-9


@pytest.mark.gpu
def test_sphere_argument_checks(bdpt):
    v, m, mats, cam, spheres = room_with_sphere_materials()
    with bdpt.Context(0) as ctx:
        with pytest.raises(bdpt.BdptError, match="radius"):
            ctx.set_spheres([((0, 0, 0), 0.0, 0)])
        with pytest.raises(bdpt.BdptError, match="256"):
            ctx.set_spheres([((0, 0, 0), 1.0, 0)] * 257)
        ctx.upload(v, m, mats)
        ctx.set_spheres([((278, 100, 278), 50.0, 99)])
        with pytest.raises(bdpt.BdptError, match="material"):
            ctx.render(bdpt.make_camera(8, 8, 1, 3, **cam), height=8)
        ctx.set_spheres(spheres)
        ctx.render(bdpt.make_camera(8, 8, 1, 3, **cam), height=8, weighting=bdpt.WEIGHT_MIS)   # spheres (and media) render under MIS too


def test_port_mis_with_analytic_spheres_agrees_with_the_reference_sum_free_estimate():
    """BDPT_WEIGHT_MIS over triangles + sphere prims (oracle/port.cpp: walk_mis asks hit_prims, visible_mis asks spheres_occlude):
    an unbiased estimator, so its mean must agree with the MIS render of the same room with the spheres finely tessellated
    (the reference-pinned triangle path), and differ from the sphere-free room."""
    from oracle.oracle import WEIGHT_MIS
    v, m, mats, cam, spheres = room_with_sphere_materials()
    W = H = 20
    spp, depth = 100, 6
    c = ocam(W, H, spp, depth, **cam)
    ana = Oracle("port").load(v, m, mats).set_spheres(spheres).render(c, rng_mode=RNG_PHILOX, seed=5, nthreads=8, height=H, weighting=WEIGHT_MIS)
    tv, tm = [np.asarray(v).reshape(-1, 9)], [np.asarray(m)]
    for ctr, rad, mat in spheres:
        tess = ts.uv_sphere(ctr, rad, 40, 80).reshape(-1, 9)
        tv.append(tess); tm.append(np.full(len(tess), mat, np.uint32))
    tes = Oracle("port").load(np.concatenate(tv), np.concatenate(tm), mats).render(c, rng_mode=RNG_PHILOX, seed=6, nthreads=8, height=H,
                                                                                    weighting=WEIGHT_MIS)
    bare = Oracle("port").load(v, m, mats).render(c, rng_mode=RNG_PHILOX, seed=5, nthreads=8, height=H, weighting=WEIGHT_MIS)
    assert abs(ana.mean() / tes.mean() - 1.0) < 0.05
    assert abs(bare.mean() / tes.mean() - 1.0) > 0.02 or np.mean(np.abs(ana - bare) > 1e-9) > 0.2
    # block-wise (5 x 5 pixel) means agree too: the spheres are where they should be
    blk = lambda f: f.reshape(H // 5, 5, W // 5, 5, 3).mean(axis=(1, 3, 4))
    assert np.abs(blk(ana) / blk(tes) - 1.0).max() < 0.35


@pytest.mark.gpu
@pytest.mark.parametrize("bvh", [0, 1])
def test_mis_sphere_film_equals_the_port_sample_for_sample(bvh, bdpt):
    from oracle.oracle import WEIGHT_MIS
    v, m, mats, cam, spheres = room_with_sphere_materials()
    W, H, spp, depth = 40, 30, 4, 8
    ref = Oracle("port").load(v, m, mats).set_spheres(spheres).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=13, nthreads=8,
                                                                       height=H, weighting=WEIGHT_MIS)
    bare = Oracle("port").load(v, m, mats).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=13, nthreads=8, height=H,
                                                  weighting=WEIGHT_MIS)
    assert rel_err(bare, ref) > 0.05
    with bdpt.Context(0) as ctx:
        ctx.upload(v, m, mats)
        ctx.set_spheres(spheres)
        ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=13, bvh=bvh, height=H, weighting=bdpt.WEIGHT_MIS)
        assert rel_err(ctx.film(), ref) < 1e-9          # fp64 atomics: only the order of the adds differs
        # together with media
        media = ts.cornell_media()
        both = Oracle("port").load(v, m, mats).set_spheres(spheres).set_media(media).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX,
                                                                                             seed=13, nthreads=8, height=H, weighting=WEIGHT_MIS)
        assert rel_err(both, ref) > 0.02
        ctx.set_media(media)
        ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=13, bvh=bvh, height=H, weighting=bdpt.WEIGHT_MIS)
        assert rel_err(ctx.film(), both) < 1e-9
        ctx.set_media([]); ctx.set_spheres([])
        ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=13, bvh=bvh, height=H, weighting=bdpt.WEIGHT_MIS)
        assert rel_err(ctx.film(), bare) < 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("bvh", [0, 1])
def test_batch_queries_ask_the_sphere_side_list(bvh, bdpt):
    """bdpt_trace / bdpt_visible with spheres set are world queries (hittable::hit on the world list, hittable.h:33): triangles,
    then the sphere prims in list order against the bound so far -- port_trace / port_visible do the same.  Bit for bit."""
    v, m, mats, cam, spheres = room_with_sphere_materials()
    port = Oracle("port").load(v, m, mats).set_spheres(spheres)
    rays = ts.ray_batch(v, 60000, seed=21, cam=cam, width=64, height=64)
    rng = np.random.default_rng(4)
    rays[::7, 7] = rng.uniform(50.0, 700.0, size=len(rays[::7]))      # some finite tmax: spheres beyond it must not answer
    p_prim, p_t, p_u, p_v = port.trace(rays, nthreads=8)
    on_sphere = (p_prim & 0xF0000000) == 0x10000000
    assert on_sphere.sum() > 3000 and (p_prim == 0xFFFFFFFF).sum() > 100 and (~on_sphere & (p_prim != 0xFFFFFFFF)).sum() > 3000
    a = rng.uniform(20.0, 530.0, size=(40000, 3)); b = rng.uniform(20.0, 530.0, size=(40000, 3))
    p_vis = port.visible(a, b, nthreads=8)
    with bdpt.Context(0) as ctx:
        ctx.upload(v, m, mats)
        bare = ctx.trace(rays, bvh=bvh)
        bare_vis = ctx.visible(a, b, bvh=bvh)
        ctx.set_spheres(spheres)
        hit = ctx.trace(rays, bvh=bvh)
        assert np.array_equal(hit["prim"], p_prim)
        assert np.array_equal(hit["t"], p_t) and np.array_equal(hit["u"], p_u) and np.array_equal(hit["v"], p_v)
        anyh = ctx.trace(rays, bvh=bvh, any_hit=True)
        assert np.array_equal(anyh["prim"] == 0xFFFFFFFF, p_prim == 0xFFFFFFFF)
        only_sphere = (bare["prim"] == 0xFFFFFFFF) & (p_prim != 0xFFFFFFFF)
        assert only_sphere.sum() > 20 and np.all((anyh["prim"][only_sphere] & 0xF0000000) == 0x10000000)
        vis = ctx.visible(a, b, bvh=bvh)
        assert np.array_equal(vis, p_vis)
        assert (bare_vis & ~vis).sum() > 500 and not (vis & ~bare_vis).any()     # spheres only ever block
        ctx.set_spheres([])
        assert np.array_equal(ctx.trace(rays, bvh=bvh)["prim"], bare["prim"])


YAML = """
camera:
  resolution: [32, 24]
  location: [278, 278, -800]
  look_at: [278, 278, 0]
  vfov: 40
  samples_per_pixel: 4
  max_depth: 8
  output: balls.png
materials:
  white: {type: lambertian, color: [186, 186, 186]}
  lamp: {type: light, emission: [15, 15, 15]}
surfaces:
  - type: mesh
    material: white
    vertices: [[0, 0, 0], [555, 0, 0], [555, 0, 555], [0, 0, 555], [0, 555, 555], [555, 555, 555]]
    triangles: [[0, 1, 2], [0, 2, 3], [3, 2, 5], [3, 5, 4]]
  - type: mesh
    material: lamp
    vertices: [[213, 554, 227], [343, 554, 227], [343, 554, 332], [213, 554, 332]]
    triangles: [[0, 1, 2], [0, 2, 3]]
  - type: Sphere
    analytic: true
    material: {type: dielectric, ior: 1.5}
    data: {center: [190, 90, 190], radius: 90}
  - type: Sphere
    analytic: true
    material: {type: metal, color: [204, 217, 224], roughness: 0.1}
    data: {center: [100, 120, 0], radius: 120}
    transform:
      rotate_y: 90
      translate: [300, 0, 250]
  - type: Sphere
    material: {type: lambertian, color: [50, 150, 75]}
    data: {center: [300, 400, 300], radius: 60}
"""


def test_yaml_analytic_flag(bdpt, tmp_path):
    """`analytic: true` makes a Sphere surface one exact prim; without it the surface is the reference's 960 triangles."""
    import math
    path = tmp_path / "balls.yaml"
    path.write_text(YAML)
    s = bdpt.SceneFile(str(path))
    assert len(s.verts) == 6 + 960 and len(s.spheres) == 2
    mats = s.material_list()
    (c0, r0, m0), (c1, r1, m1) = s.spheres
    assert c0 == (190.0, 90.0, 190.0) and r0 == 90.0 and mats[m0][0] == bdpt.MAT_DIELECTRIC and mats[m0][2] == 1.5
    r = 90 * 3.1415926535897932385 / 180.0
    c, sn = math.cos(r), math.sin(r)
    assert c1 == ((c * 100.0) + (sn * 0.0) + 300, 120.0, (-sn * 100.0) + (c * 0.0) + 250) and r1 == 120.0   # rotate_y, then translate
    assert mats[m1][0] == bdpt.MAT_METAL and mats[m1][2] == 0.1 and m0 != m1 and max(m0, m1) == len(mats) - 1
    bad = tmp_path / "bad.yaml"
    bad.write_text(YAML.replace("material: {type: dielectric, ior: 1.5}", "material: {type: light, emission: [5, 5, 5]}"))
    with pytest.raises(bdpt.BdptError, match="light"):
        bdpt.SceneFile(str(bad))


@pytest.mark.gpu
def test_cli_renders_a_yaml_scene_with_analytic_spheres(bdpt, tmp_path):
    import os
    import subprocess
    path = tmp_path / "balls.yaml"
    path.write_text(YAML)
    exe = os.path.join(os.path.dirname(bdpt.LIB_PATH), "bdpt")
    r = subprocess.run([exe, str(path), "--seed", "4", "--quiet"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    s = bdpt.SceneFile(str(path))
    with bdpt.Context(0) as ctx:
        ctx.upload(s.verts, s.mat_id, s.materials)
        ctx.set_spheres(s.spheres)
        ctx.render(s.camera, seed=4, bvh=bdpt.BVH_FAST)
        rgb = ctx.film_rgb8(4)
        ctx.set_spheres([])
        ctx.render(s.camera, seed=4, bvh=bdpt.BVH_FAST)
        bare = ctx.film_rgb8(4)
    assert (rgb != bare).mean() > 0.05
    ref_png = tmp_path / "ctypes.png"
    bdpt.write_png(str(ref_png), rgb)
    assert (tmp_path / "output" / "balls.png").read_bytes() == ref_png.read_bytes()
    r = subprocess.run([exe, str(path), "--cache", str(tmp_path / "c.bdptimg"), "--quiet"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 1 and "without --cache" in r.stderr
