"""The fast tree built on the device (bdpt_scene_upload_ex + BDPT_UPLOAD_FAST_TREE_GPU, csrc/gpu_bvh.cu)
must change nothing but speed: same prim IDs, t, u, v as the reference golden vectors and the oracle port, the
same visibility booleans, the same film as the host-built tree, and a depth the walk's stack can hold."""
import numpy as np
import pytest

from oracle.oracle import Oracle
from test_trace_gpu import NAMES, check_hits

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    return float(np.max(np.abs(a - b)) / max(1e-300, float(np.max(np.abs(b)))))


@pytest.mark.parametrize("name", NAMES)
def test_gpu_tree_trace_matches_reference_golden(name, bdpt, gpu_ctx, scenes, golden):
    g = golden(name)
    v, m, mats, cam = scenes[name]()
    gpu_ctx.upload(v, m, mats, gpu_tree=True)
    n_nodes, depth = gpu_ctx.bvh_info(bdpt.BVH_FAST)
    assert 1 <= n_nodes <= len(v) and 2 <= depth <= 30
    assert gpu_ctx.bvh_info(bdpt.BVH_REFERENCE)[0] == int(g["nodes"])
    h = gpu_ctx.trace(g["rays"], bvh=bdpt.BVH_FAST)
    check_hits(h, g["prim"], g["t"], g["u"], g["v"], exact=True)
    vis = gpu_ctx.visible(g["seg_a"], g["seg_b"], bvh=bdpt.BVH_FAST)
    assert np.array_equal(vis, g["visible"])


def test_gpu_tree_large_soup_vs_port(bdpt, gpu_ctx):
    from tools import testscenes as ts
    v, m, mats = ts.random_soup(50000, seed=9, extent=20.0, size=0.7)
    rays = ts.ray_batch(v, 300000, seed=123)
    o = Oracle("port").load(v, m, mats)
    prim, t, u, w = o.trace(rays, nthreads=8)
    gpu_ctx.upload(v, m, mats, gpu_tree=True)
    h = gpu_ctx.trace(rays, bvh=bdpt.BVH_FAST)
    check_hits(h, prim, t, u, w, exact=True)
    ha = gpu_ctx.trace(rays, bvh=bdpt.BVH_FAST, any_hit=True)
    assert np.array_equal(ha["prim"] != 0xFFFFFFFF, prim != 0xFFFFFFFF)


@pytest.mark.parametrize("n", [1, 2, 3, 5])
def test_gpu_tree_tiny_scenes(n, bdpt, gpu_ctx):
    """one- and two-triangle scenes (the root needs two children), and identical triangles (equal Morton codes)"""
    rng = np.random.default_rng(n)
    tri = np.array([[-1, -1, 0, 1, -1, 0, 0, 1, 0]], np.float64)
    v = np.repeat(tri, n, axis=0)
    v[:, 2::3] += np.arange(n)[:, None] * (0.0 if n == 5 else 0.5)    # n == 5: five coincident triangles
    m = np.zeros(n, np.uint32)
    mats = [(bdpt.MAT_LAMBERTIAN, (0.5, 0.5, 0.5), 0.0)]
    rays = np.zeros((64, 8))
    rays[:, 0:2] = rng.uniform(-1.2, 1.2, (64, 2)); rays[:, 2] = -3.0
    rays[:, 5] = 1.0; rays[:, 6] = 0.001; rays[:, 7] = np.inf
    gpu_ctx.upload(v, m, mats)
    want = gpu_ctx.trace(rays, bvh=bdpt.BVH_REFERENCE)
    gpu_ctx.upload(v, m, mats, gpu_tree=True)
    got = gpu_ctx.trace(rays, bvh=bdpt.BVH_FAST)
    check_hits(got, want["prim"], want["t"], want["u"], want["v"], exact=True)


def test_gpu_tree_depth_is_bounded_on_clustered_input(bdpt, gpu_ctx):
    """200k slivers crowded into a 2^-30 corner of the bounds plus one far triangle: a radix tree over the codes
    would be as deep as the codes are long; the builder's clamp keeps the walk's 32-slot stack safe."""
    rng = np.random.default_rng(5)
    n = 200000
    base = rng.uniform(0.0, 1e-6, (n, 1, 3))
    v = (base + rng.uniform(0.0, 1e-7, (n, 3, 3))).reshape(n, 9)
    v = np.vstack([v, np.array([[1000, 1000, 1000, 1001, 1000, 1000, 1000, 1001, 1000]], np.float64)])
    m = np.zeros(len(v), np.uint32)
    mats = [(bdpt.MAT_LAMBERTIAN, (0.5, 0.5, 0.5), 0.0)]
    gpu_ctx.upload(v, m, mats, gpu_tree=True)
    n_nodes, depth = gpu_ctx.bvh_info(bdpt.BVH_FAST)
    assert depth <= 30 and n_nodes < len(v)
    from tools import testscenes as ts
    rays = ts.ray_batch(v[:n], 20000, seed=3)
    got = gpu_ctx.trace(rays, bvh=bdpt.BVH_FAST)
    want = gpu_ctx.trace(rays, bvh=bdpt.BVH_REFERENCE)
    assert np.array_equal(got["prim"], want["prim"]) and np.array_equal(got["t"], want["t"])


def test_gpu_tree_renders_the_same_film(bdpt, gpu_ctx, scenes):
    v, m, mats, cam = scenes["cornell_spheres"]()
    c = bdpt.make_camera(24, 24, 16, 8, **cam)
    gpu_ctx.upload(v, m, mats)
    gpu_ctx.render(c, seed=11, height=24, bvh=bdpt.BVH_FAST)
    want = gpu_ctx.film()
    gpu_ctx.upload(v, m, mats, gpu_tree=True)
    gpu_ctx.render(c, seed=11, height=24, bvh=bdpt.BVH_FAST)
    assert rel_err(gpu_ctx.film(), want) < 1e-12   # fp64 atomics: summation order only
