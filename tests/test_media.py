"""Participating media: constant_medium + isotropic (/root/reference/src/materials/volumes/constant_medium.h:8-61,
materials/material.h:149-171) -- bdpt_scene_set_media.
  CPU: the port's restatement (oracle/port.cpp: hit_world / medium_hit / visible_media) against the reference's OWN
       constant_medium objects in a bvh_node world (oracle/_ref): the two draw their random numbers in different orders
       (the reference whenever its BVH walk reaches the medium, the port after the triangle walk), so the pin is
       statistical -- image means within Monte-Carlo error, and far from the medium-free image.
  GPU: film equal to the port's sample for sample on the counter-based stream (tolerance 1e-9 of the peak: fp64 atomics
       reorder the adds), both tree flavours; the draw inside constant_medium::hit goes through a deterministic log shared
       bit for bit by the kernels and the port.
  BDPT_WEIGHT_MIS with media (volume vertices in the recursive balance-heuristic weights, transmittance in every pdf and
  contribution): the port's estimator against the reference's OWN path tracer (camera::path_trace_color) run over the same
  constant_medium objects -- equal means, an unbiased estimator; GPU film = port film sample for sample."""
import numpy as np
import pytest

from oracle.oracle import Oracle, available, make_camera as ocam, RNG_MT19937, RNG_PHILOX
from tools import testscenes as ts


def rel_err(a, b):
    return float(np.abs(a - b).max() / max(1e-300, np.abs(b).max()))


@pytest.mark.skipif(not available("ref"), reason="oracle/_ref not shipped")
def test_port_media_agree_with_the_references_own_constant_medium():
    v, m, mats, cam = ts.cornell(False)
    media = ts.cornell_media()
    W = H = 20
    spp, depth, seeds = 36, 6, 4
    c = ocam(W, H, spp, depth, **cam)
    port = sum(Oracle("port").load(v, m, mats).set_media(media).render(c, rng_mode=RNG_MT19937, seed=s, height=H) for s in range(1, 1 + seeds))
    ref = sum(Oracle("ref").load(v, m, mats).set_media(media).render(c, rng_mode=RNG_MT19937, seed=50 + s, height=H) for s in range(1, 1 + seeds))
    clear = sum(Oracle("port").load(v, m, mats).render(c, rng_mode=RNG_MT19937, seed=s, height=H) for s in range(1, 1 + seeds))
    mp, mr = port.mean(axis=(0, 1)), ref.mean(axis=(0, 1))
    assert np.all(np.abs(mp / mr - 1.0) < 0.06), (mp, mr)           # 4 x 36 spp x 400 pixels: ~2 % noise on the mean
    assert clear.mean() / ref.mean() > 1.08                          # the smoke is not a no-op (per-seed means: 7.8 against 6.8)


def test_det_log_is_log_to_the_last_bit_or_so():
    import ctypes as C
    import math
    f = Oracle("port").lib.port_det_log
    f.restype = C.c_double; f.argtypes = [C.c_double]
    rng = np.random.default_rng(5)
    xs = np.concatenate([rng.random(20000), rng.random(500) * 1e-12, [2.0 ** -53, 0.5, 1 - 2.0 ** -53, 0.70710678, 0.70710679]])
    assert max(abs(f(float(x)) - math.log(x)) / abs(math.log(x)) for x in xs) < 4e-16
    assert f(0.0) == -math.inf


@pytest.mark.gpu
@pytest.mark.parametrize("bvh", [0, 1])
@pytest.mark.parametrize("name", ["cornell_box", "cornell_spheres"])
def test_media_film_equals_the_port_sample_for_sample(name, bvh, bdpt, scenes):
    v, m, mats, cam = scenes[name]()
    media = ts.cornell_media()
    W, H, spp, depth = 36, 28, 4, 7
    ref = Oracle("port").load(v, m, mats).set_media(media).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=42,
                                                                  nthreads=8, height=H)
    clear = Oracle("port").load(v, m, mats).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=42, nthreads=8, height=H)
    assert rel_err(clear, ref) > 0.05                              # the media change the image
    with bdpt.Context(0) as ctx:
        ctx.set_media(media)                                        # before the upload: belongs to the context
        ctx.upload(v, m, mats)
        ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=42, bvh=bvh, height=H)
        assert rel_err(ctx.film(), ref) < 1e-9
        ctx.set_media([])                                           # dropped: the plain scene again
        ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=42, bvh=bvh, height=H)
        assert rel_err(ctx.film(), clear) < 1e-9


@pytest.mark.gpu
def test_media_argument_checks(bdpt):
    with bdpt.Context(0) as ctx:
        box = ts.box_mesh((0, 0, 0), (1, 1, 1))
        with pytest.raises(bdpt.BdptError, match="density"):
            ctx.set_media([(0.0, (1, 1, 1), box)])
        with pytest.raises(bdpt.BdptError, match="16"):
            ctx.set_media([(1.0, (1, 1, 1), box)] * 17)


YAML = """
camera:
  resolution: [24, 24]
  location: [278, 278, -800]
  look_at: [278, 278, 0]
  vfov: 40
  samples_per_pixel: 4
  max_depth: 6
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
media:
  - density: 0.01
    color: [230, 230, 230]
    box: [[100, 0, 100], [300, 300, 300]]
  - density: 0.004
    color: [0.2, 0.4, 0.9]
    vertices: [[0, 0, 0], [250, 0, 0], [0, 250, 0], [0, 0, 200]]
    triangles: [[0, 2, 1], [0, 1, 3], [0, 3, 2], [1, 2, 3]]
    transform:
      rotate_y: 30
      translate: [250, 150, 50]
"""


def test_yaml_media_section(bdpt, tmp_path):
    """The loader's `media:` section (an extension, like `texture:` and `transform:`): a box and a transformed tetrahedron."""
    import math
    path = tmp_path / "smoke.yaml"
    path.write_text(YAML)
    s = bdpt.SceneFile(str(path))
    assert len(s.media) == 2 and len(s.verts) == 6
    d0, c0, b0 = s.media[0]
    assert d0 == 0.01 and np.allclose(c0, [230 / 255.0] * 3, rtol=0, atol=1e-15) and b0.shape == (12, 9)
    assert np.array_equal(b0.reshape(-1, 3).min(axis=0), [100, 0, 100]) and np.array_equal(b0.reshape(-1, 3).max(axis=0), [300, 300, 300])
    assert np.array_equal(np.sort(b0.reshape(-1, 9), axis=0), np.sort(ts.box_mesh((100, 0, 100), (300, 300, 300)), axis=0))
    d1, c1, b1 = s.media[1]
    assert d1 == 0.004 and c1 == (0.2, 0.4, 0.9) and b1.shape == (4, 9)
    r = 30 * 3.1415926535897932385 / 180.0
    c, sn = math.cos(r), math.sin(r)
    corner = np.array([250.0, 0.0, 0.0])                          # vertex 1 of the tetrahedron: rotate_y, then translate
    want = np.array([(c * corner[0]) + (sn * corner[2]) + 250, corner[1] + 150, (-sn * corner[0]) + (c * corner[2]) + 50])
    assert any(np.array_equal(p, want) for p in b1.reshape(-1, 3))
    bad = tmp_path / "bad.yaml"
    bad.write_text(YAML.replace("density: 0.01", "density: 0"))
    with pytest.raises(bdpt.BdptError, match="density"):
        bdpt.SceneFile(str(bad))


@pytest.mark.gpu
def test_cli_renders_a_yaml_scene_with_media(bdpt, tmp_path):
    """bdpt scene.yaml (host/main.cpp -> camera::render -> bdpt_scene_set_media) writes the PNG the ctypes path computes."""
    import os
    import subprocess
    path = tmp_path / "smoke.yaml"
    path.write_text(YAML.replace("resolution: [24, 24]", "resolution: [24, 24]\n  output: smoke.png"))
    exe = os.path.join(os.path.dirname(bdpt.LIB_PATH), "bdpt")
    r = subprocess.run([exe, str(path), "--seed", "9", "--quiet"], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    s = bdpt.SceneFile(str(path))
    with bdpt.Context(0) as ctx:
        ctx.set_media(s.media)
        ctx.upload(s.verts, s.mat_id, s.materials)
        ctx.render(s.camera, seed=9, bvh=bdpt.BVH_FAST)
        rgb = ctx.film_rgb8(4)
        ctx.set_media([])
        ctx.render(s.camera, seed=9, bvh=bdpt.BVH_FAST)
        clear = ctx.film_rgb8(4)
    assert (rgb != clear).mean() > 0.05
    ref_png = tmp_path / "ctypes.png"
    bdpt.write_png(str(ref_png), rgb)
    assert (tmp_path / "output" / "smoke.png").read_bytes() == ref_png.read_bytes()
    # --mis: render_options.weighting = BDPT_WEIGHT_MIS through the same front end (media under MIS: volume vertices)
    r = subprocess.run([exe, str(path), "--seed", "9", "--quiet", "--mis", "--out", "smoke_mis.png"], cwd=tmp_path, capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stderr
    with bdpt.Context(0) as ctx:
        ctx.set_media(s.media)
        ctx.upload(s.verts, s.mat_id, s.materials)
        ctx.render(s.camera, seed=9, bvh=bdpt.BVH_FAST, weighting=bdpt.WEIGHT_MIS)
        mis_rgb = ctx.film_rgb8(4)
    assert (mis_rgb != rgb).mean() > 0.05
    mis_png = tmp_path / "ctypes_mis.png"
    bdpt.write_png(str(mis_png), mis_rgb)   # (fp64 atomics reorder the adds by ~1e-16: far from an 8-bit step)
    assert (tmp_path / "output" / "smoke_mis.png").read_bytes() == mis_png.read_bytes()


@pytest.mark.skipif(not available("ref"), reason="oracle/_ref not shipped")
def test_port_mis_with_media_agrees_with_the_references_path_tracer():
    """The MIS estimator with volume vertices is unbiased: same image mean as camera::path_trace_color (camera.h:255-292) over the
    reference's own constant_medium objects, overall and per 4 x 4-pixel block; the medium-free MIS image is 12 % brighter."""
    from oracle.oracle import WEIGHT_MIS
    v, m, mats, cam = ts.cornell(False)
    media = ts.cornell_media()
    W = H = 16
    c = ocam(W, H, 900, 6, **cam)
    mis = Oracle("port").load(v, m, mats).set_media(media).render(c, rng_mode=RNG_PHILOX, seed=3, nthreads=8, height=H, weighting=WEIGHT_MIS)
    pt = Oracle("ref").load(v, m, mats).set_media(media).render_pt(c, seed=5, height=H)
    clear = Oracle("port").load(v, m, mats).render(c, rng_mode=RNG_PHILOX, seed=3, nthreads=8, height=H, weighting=WEIGHT_MIS)
    assert abs(mis.mean() / pt.mean() - 1.0) < 0.02, (mis.mean(), pt.mean())
    assert clear.mean() / pt.mean() > 1.08
    blk = lambda f: f.reshape(4, 4, 4, 4, 3).mean(axis=(1, 3, 4))
    assert np.abs(blk(mis) / blk(pt) - 1.0).max() < 0.12


@pytest.mark.gpu
@pytest.mark.parametrize("bvh", [0, 1])
@pytest.mark.parametrize("name", ["cornell_box", "cornell_spheres"])
def test_mis_media_film_equals_the_port_sample_for_sample(name, bvh, bdpt, scenes):
    from oracle.oracle import WEIGHT_MIS
    v, m, mats, cam = scenes[name]()
    media = ts.cornell_media()
    W, H, spp, depth = 36, 28, 4, 7
    ref = Oracle("port").load(v, m, mats).set_media(media).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=42,
                                                                  nthreads=8, height=H, weighting=WEIGHT_MIS)
    clear = Oracle("port").load(v, m, mats).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=42, nthreads=8, height=H,
                                                   weighting=WEIGHT_MIS)
    assert np.mean(np.abs(clear - ref) > 1e-6 * ref.max()) > 0.5     # the media change the image (nearly every pixel)
    with bdpt.Context(0) as ctx:
        ctx.set_media(media)
        ctx.upload(v, m, mats)
        ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=42, bvh=bvh, height=H, weighting=bdpt.WEIGHT_MIS)
        assert rel_err(ctx.film(), ref) < 1e-9          # fp64 atomics reorder the adds; exp() of the transmittance: an ulp
        ctx.set_media([])
        ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=42, bvh=bvh, height=H, weighting=bdpt.WEIGHT_MIS)
        assert rel_err(ctx.film(), clear) < 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("weighting", [0, 1])
def test_large_boundary_goes_through_its_tree_and_gives_the_port_film(weighting, bdpt):
    """A medium bounded by a tessellated sphere (528 triangles) beside a box-shaped one: bdpt_scene_set_media builds a tree over the
    large boundary (more than 16 triangles); the port loops over all of them.  Same film, sample for sample."""
    v, m, mats, cam = ts.cornell(False)
    ball = ts.uv_sphere((278.0, 200.0, 278.0), 120.0, 12, 24).reshape(-1, 9)
    assert len(ball) > 400
    media = [(0.012, (0.8, 0.85, 0.9), ball), (0.004, (0.2, 0.4, 0.9), ts.box_mesh((250, 150, 50), (500, 400, 250)))]
    W, H, spp, depth = 32, 24, 4, 6
    ref = Oracle("port").load(v, m, mats).set_media(media).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=23, nthreads=8,
                                                                  height=H, weighting=weighting)
    clear = Oracle("port").load(v, m, mats).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=23, nthreads=8, height=H,
                                                   weighting=weighting)
    assert np.mean(np.abs(clear - ref) > 1e-6 * ref.max()) > 0.5
    for bvh in (0, 1):
        with bdpt.Context(0) as ctx:
            ctx.upload(v, m, mats)
            ctx.set_media(media)
            ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=23, bvh=bvh, height=H, weighting=weighting)
            assert rel_err(ctx.film(), ref) < 1e-9
