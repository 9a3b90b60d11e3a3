"""GPU parity of the wavefront BDPT integrator (bdpt_render) against the CPU oracle, sample for sample.

The oracle port (oracle/port.cpp) is bit-identical to the reference's own headers in its mt19937 mode
(tests/test_oracle.py, tests/golden/).  Its counter-based mode changes only the random stream (and uses
libm-independent sin/cos), and the CUDA kernels use that same stream and the same fp64 operation
order, so every sample's radiance must agree; pixel sums differ only by fp64 summation order.
Tolerance: 1e-9 of the image's peak value (fp64 atomics reorder at most ~1e3 addends per pixel)."""
import numpy as np
import pytest

from oracle.oracle import Oracle, make_camera as ocam, RNG_PHILOX

pytestmark = pytest.mark.gpu

CASES = [  # scene, W, H, spp, depth
    ("cornell_box", 24, 24, 4, 5),
    ("cornell_spheres", 32, 32, 4, 8),
    ("unit_room", 32, 20, 9, 12),
    ("cornell_spheres", 16, 16, 16, 24),
    ("sphere_grid", 32, 18, 4, 32),   # mixed dielectric / fuzzy-metal spheres, depth 32: the divergence-stress config in miniature
]


def rel_err(a, b):
    return np.abs(a - b).max() / max(1e-300, np.abs(b).max())


@pytest.mark.parametrize("name,W,H,spp,depth", CASES)
@pytest.mark.parametrize("bvh", [0, 1])
def test_render_matches_oracle_per_sample(name, W, H, spp, depth, bvh, bdpt, gpu_ctx, scenes):
    v, m, mats, cam = scenes[name]()
    ref = Oracle("port").load(v, m, mats).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=42,
                                                 nthreads=8, height=H)
    gpu_ctx.upload(v, m, mats)
    gpu_ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=42, bvh=bvh, height=H)
    film = gpu_ctx.film()
    assert film.shape == ref.shape
    assert rel_err(film, ref) < 1e-9
    st = gpu_ctx.stats()
    assert st["kernel_launches"] > 0 and st["rays_closest"] > 0


def test_render_is_independent_of_batching_and_sharding(bdpt, gpu_ctx, scenes):
    """strata sharded over ranks and small wavefront batches give the same film."""
    v, m, mats, cam = scenes["cornell_spheres"]()
    gpu_ctx.upload(v, m, mats)
    c = bdpt.make_camera(20, 20, 16, 8, **cam)
    gpu_ctx.render(c, seed=3, height=20)
    whole = gpu_ctx.film()
    gpu_ctx.film_clear(20, 20)
    for g in range(4):   # 4 "ranks", strata g, g+4, ...
        gpu_ctx.render(c, seed=3, strata=(g, 0, 4), batch_samples=777, clear=False, height=20)
    assert rel_err(gpu_ctx.film(), whole) < 1e-12
    gpu_ctx.film_clear(20, 20)
    gpu_ctx.render(c, seed=3, rows=(0, 7), clear=False, height=20)
    gpu_ctx.render(c, seed=3, rows=(7, 20), clear=False, height=20)
    assert rel_err(gpu_ctx.film(), whole) < 1e-12


def test_render_rays_and_vertices_match_oracle_counts(bdpt, gpu_ctx, scenes):
    v, m, mats, cam = scenes["cornell_box"]()
    o = Oracle("port").load(v, m, mats)
    o.stats_reset()
    o.render(ocam(16, 16, 4, 6, **cam), rng_mode=RNG_PHILOX, seed=9, height=16)
    s = o.stats()
    gpu_ctx.upload(v, m, mats)
    gpu_ctx.stats_reset()
    gpu_ctx.render(bdpt.make_camera(16, 16, 4, 6, **cam), seed=9, height=16)
    g = gpu_ctx.stats()
    assert g["samples"] == 16 * 16 * 4
    assert g["rays_closest"] == s["rays"] - s["shadow_rays"]
    # the kernels skip connections whose contribution is exactly zero before tracing them
    assert 0 < g["rays_shadow"] <= s["shadow_rays"]


def test_tonemap_matches_oracle(bdpt, gpu_ctx, scenes):
    v, m, mats, cam = scenes["cornell_box"]()
    gpu_ctx.upload(v, m, mats)
    gpu_ctx.render(bdpt.make_camera(16, 16, 4, 5, **cam), seed=1, height=16)
    film = gpu_ctx.film()
    assert np.array_equal(gpu_ctx.film_rgb8(4), Oracle("port").colors_to_rgb8(film, 4))


def test_render_edge_cases(bdpt, gpu_ctx, scenes):
    v, m, mats, cam = scenes["cornell_box"]()
    gpu_ctx.upload(v, m, mats)
    gpu_ctx.render(bdpt.make_camera(8, 8, 1, 0, **cam), height=8)      # depth 0 -> black (camera.h:333)
    assert not gpu_ctx.film().any()
    gpu_ctx.render(bdpt.make_camera(8, 8, 3, 3, **cam), height=8)      # spp 3 -> sqrt_spp 1 (camera.h:164)
    assert gpu_ctx.stats()["samples"] >= 64
    # no lights: every triangle is in the light list but emits nothing -> only background (main.cpp:67)
    nolight = [(t if t != 3 else 0, rgb, p) for t, rgb, p in mats]
    gpu_ctx.upload(v, m, nolight)
    c = bdpt.make_camera(8, 8, 1, 4, lookfrom=cam["lookfrom"], lookat=cam["lookat"], vfov=cam["vfov"], background=(0.5, 0.25, 1.0))
    gpu_ctx.render(c, height=8)
    ref = Oracle("port").load(v, m, nolight).render(
        ocam(8, 8, 1, 4, lookfrom=cam["lookfrom"], lookat=cam["lookat"], vfov=cam["vfov"], background=(0.5, 0.25, 1.0)),
        rng_mode=RNG_PHILOX, seed=0, height=8)
    assert rel_err(gpu_ctx.film(), ref) < 1e-9 or not ref.any()


def test_split_walk_and_parking_keep_the_film(bdpt, gpu_ctx, scenes, monkeypatch):
    """BDPT_WORK_LIMIT parks every ray that touches more than n nodes + triangles and re-runs it split 32 ways over
    a warp (traverse_walk.cuh: split_walk); the film must not move."""
    v, m, mats, cam = scenes["cornell_spheres"]()
    gpu_ctx.upload(v, m, mats)
    c = bdpt.make_camera(24, 24, 9, 10, **cam)
    monkeypatch.delenv("BDPT_WORK_LIMIT", raising=False)
    gpu_ctx.render(c, seed=11, bvh=bdpt.BVH_FAST, height=24)
    whole = gpu_ctx.film()
    for limit in ("8", "40"):
        monkeypatch.setenv("BDPT_WORK_LIMIT", limit)
        gpu_ctx.stats_reset()
        gpu_ctx.render(c, seed=11, bvh=bdpt.BVH_FAST, height=24, batch_samples=1500)
        assert rel_err(gpu_ctx.film(), whole) < 1e-12
        st = gpu_ctx.stats()
        assert st["deferred_closest"] > 0 and st["deferred_shadow"] > 0   # the parked paths really ran


@pytest.mark.parametrize("name,W,H,spp,depth", [("coffee_machine", 96, 54, 4, 16), ("glass-of-water", 64, 64, 4, 80)])
def test_named_scenes_match_oracle_per_sample(name, W, H, spp, depth, bdpt, gpu_ctx):
    """BASELINE.json's named scenes (stand-in geometry: the reference ships no OBJ assets), loaded through
    the YAML front-end, camera from the file: GPU film == oracle port film, sample for sample, on both BVH flavours."""
    import os
    from tools import make_standin_scenes as ms
    ms.generate(detail=2, quiet=True)
    sf = bdpt.SceneFile(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes", name + ".yaml"))
    cam = sf.camera_with(W, H, spp, depth)
    ref = Oracle("port").load(sf.verts, sf.mat_id, sf.materials).render(cam, rng_mode=RNG_PHILOX, seed=5, nthreads=16, height=H)
    assert np.abs(ref).max() > 0
    gpu_ctx.upload(sf.verts, sf.mat_id, sf.materials)
    for bvh in (bdpt.BVH_FAST, bdpt.BVH_REFERENCE):
        gpu_ctx.render(cam, seed=5, bvh=bvh, height=H)
        assert rel_err(gpu_ctx.film(), ref) < 1e-9


def test_cli_renders_the_same_png_as_the_api(bdpt, gpu_ctx, tmp_path):
    """`bdpt scene.yaml` (host/main.cpp = the reference's main.cpp:62-78 on libbdpt): load_scene_from_yaml ->
    bvh_node -> camera::render -> ./output/<name>.  Its PNG must be the C-ABI path's film through wpng.h's
    colors_to_rgb8, byte for byte."""
    import os, subprocess
    from test_loader import decode_png
    from tools import make_standin_scenes as ms
    ms.generate(detail=2, quiet=True)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "bidirectional-path-tracer_b200", "bdpt")
    scene = os.path.join(root, "scenes", "coffee_machine.yaml")
    r = subprocess.run([exe, scene, "--res", "64x36", "--spp", "4", "--depth", "6", "--seed", "3", "--out", "cli.png"],
                       cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-1000:]
    assert "Triangles:" in r.stdout                                    # scene_loader.h:521
    png = decode_png(tmp_path / "output" / "cli.png")                  # wpng.h:45-49: ./output/<file_name>
    sf = bdpt.SceneFile(scene)
    cam = sf.camera_with(64, 36, 4, 6)
    gpu_ctx.upload(sf.verts, sf.mat_id, sf.materials)
    gpu_ctx.render(cam, seed=3, bvh=bdpt.BVH_FAST, height=36)
    rgb = gpu_ctx.film_rgb8(4)
    assert png.shape == rgb.shape == (36, 64, 3)
    # fp64 atomics sum a pixel's samples in a different order in the two runs: allow one 8-bit step on rare pixels
    d = np.abs(png.astype(int) - rgb.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 0.01
    # a bad argument fails like main.cpp:72-75
    r = subprocess.run([exe, str(tmp_path / "missing.yaml")], cwd=tmp_path, capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "Failed to load scene" in r.stderr
