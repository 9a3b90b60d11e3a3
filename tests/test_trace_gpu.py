"""GPU parity of the BVH traversal kernels (k_trace / k_visible) through the C-ABI, against the oracle
port and the committed reference golden vectors. Prim IDs must be bit-exact; t within 1e-5 relative
(BASELINE.json north_star) — the reference-topology kernel is in fact bit-exact in t, u, v too."""
import numpy as np
import pytest

from oracle.oracle import Oracle

pytestmark = pytest.mark.gpu
NAMES = ["cornell_spheres", "cornell_box", "unit_room"]


def check_hits(h, prim, t, u, v, exact):
    assert np.array_equal(h["prim"], prim)
    hit = prim != 0xFFFFFFFF
    if exact:
        assert np.array_equal(h["t"], t) and np.array_equal(h["u"], u) and np.array_equal(h["v"], v)
    else:
        assert np.all(np.abs(h["t"][hit] - t[hit]) <= 1e-5 * np.abs(t[hit]))   # tolerance: 1e-5 relative
    assert np.all(h["t"][~hit] == 0.0)


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("bvh", [0, 1])
def test_trace_matches_reference_golden(name, bvh, bdpt, gpu_ctx, scenes, golden):
    g = golden(name)
    v, m, mats, cam = scenes[name]()
    gpu_ctx.upload(v, m, mats)
    assert gpu_ctx.bvh_info(0)[0] == int(g["nodes"])
    assert np.array_equal(gpu_ctx.leaf_order(), g["leaf_order"])
    h = gpu_ctx.trace(g["rays"], bvh=bvh)
    check_hits(h, g["prim"], g["t"], g["u"], g["v"], exact=True)
    vis = gpu_ctx.visible(g["seg_a"], g["seg_b"], bvh=bvh)
    assert np.array_equal(vis, g["visible"])


@pytest.mark.parametrize("bvh", [0, 1])
def test_trace_large_soup_vs_port(bvh, bdpt, gpu_ctx):
    from tools import testscenes as ts
    v, m, mats = ts.random_soup(50000, seed=9, extent=20.0, size=0.7)
    rays = ts.ray_batch(v, 300000, seed=123)
    o = Oracle("port").load(v, m, mats)
    prim, t, u, w = o.trace(rays, nthreads=8)
    gpu_ctx.upload(v, m, mats)
    h = gpu_ctx.trace(rays, bvh=bvh)
    check_hits(h, prim, t, u, w, exact=True)
    # any-hit: same boolean as closest-hit
    ha = gpu_ctx.trace(rays, bvh=bvh, any_hit=True)
    assert np.array_equal(ha["prim"] != 0xFFFFFFFF, prim != 0xFFFFFFFF)


def test_trace_stats_counts_match_oracle(bdpt, gpu_ctx, scenes):
    """Reference-topology kernel visits exactly the nodes the reference visits (stats.h counters)."""
    v, m, mats, cam = scenes["cornell_spheres"]()
    from tools import testscenes as ts
    rays = ts.ray_batch(v, 20000, seed=5, cam=cam)
    o = Oracle("port").load(v, m, mats)
    o.stats_reset()
    o.trace(rays)
    s = o.stats()
    gpu_ctx.upload(v, m, mats)
    gpu_ctx.stats_reset()
    gpu_ctx.trace(rays, bvh=0, stats=True)
    gs = gpu_ctx.stats()
    assert gs["node_visits"] == s["node_visits"]
    assert gs["tri_tests"] == s["tri_tests"]
    # the rest of the reference's BvhStats (stats.h:8-16): every counter of its end-of-render block
    assert gs["aabb_tests"] == s["aabb_tests"] == s["node_visits"]
    assert gs["aabb_hits"] == s["aabb_hits"]
    assert gs["triangle_hits"] == s["tri_hits"]
    assert gs["bvh_nodes_built"] == o.node_count()
    gpu_ctx.stats_reset()
    gpu_ctx.trace(rays, bvh=1, stats=True)
    fs = gpu_ctx.stats()
    assert 0 < fs["node_visits"] < s["node_visits"]   # the fast tree must not visit more than the reference


def test_trace_edge_cases(bdpt, gpu_ctx):
    """Empty batch, rays with zero direction components (inf / NaN slabs), shared-edge ties, t at the
    inclusive interval ends, a single-triangle scene, duplicated triangles."""
    from tools import testscenes as ts
    q = ts.quad((0, 0, 0), (1, 0, 0), (0, 1, 0))            # two triangles sharing the diagonal
    v = np.concatenate([q, q, ts.quad((0, 0, 1), (1, 0, 0), (0, 1, 0))])
    m = np.zeros(len(v), np.uint32)
    mats = [(0, (0.5, 0.5, 0.5), 0.0)]
    rays = []
    for x, y in [(0.5, 0.5), (0.25, 0.75), (0.75, 0.25), (0.0, 0.0), (1.0, 1.0), (0.5, 0.0), (0.3, 0.3)]:
        rays.append([x, y, -1.0, 0, 0, 1.0, 0.001, np.inf])      # axis-parallel: two zero components
        rays.append([x, y, -1.0, 0, 0, 2.0, 0.001, 0.5])         # tmax exactly at the hit (t = 0.5)
        rays.append([x, y, -1.0, 0, 0, 1.0, 1.0, 5.0])           # tmin exactly at the hit
        rays.append([x, y, 2.0, 0, 0, -1.0, 0.001, np.inf])      # from the other side
    rays.append([0.5, 0.5, -1, 1, 0, 0, 0.001, np.inf])          # parallel to the quads: miss
    rays = np.array(rays, dtype=np.float64)
    o = Oracle("port").load(v, m, mats)
    prim, t, u, w = o.trace(rays)
    gpu_ctx.upload(v, m, mats)
    assert len(gpu_ctx.trace(np.zeros((0, 8)))) == 0
    for bvh in (0, 1):
        check_hits(gpu_ctx.trace(rays, bvh=bvh), prim, t, u, w, exact=True)
    # single triangle
    v1 = v[:1]
    o1 = Oracle("port").load(v1, m[:1], mats)
    r1 = rays[:8]
    gpu_ctx.upload(v1, m[:1], mats)
    for bvh in (0, 1):
        check_hits(gpu_ctx.trace(r1, bvh=bvh), *o1.trace(r1), exact=True)


def test_state_errors(bdpt):
    ctx = bdpt.Context(0)
    with pytest.raises(bdpt.BdptError, match="no scene"):
        ctx.trace(np.zeros((1, 8)))
    ctx.close()


@pytest.mark.parametrize("name", NAMES)
def test_split_walk_matches_reference_golden(name, bdpt, gpu_ctx, scenes, golden):
    """The warp-cooperative walk (one ray split 32 ways along its length, traverse_walk.cuh) returns the reference's
    hits bit for bit: it is the path bdpt_render sends rays over its work limit through."""
    g = golden(name)
    v, m, mats, cam = scenes[name]()
    gpu_ctx.upload(v, m, mats)
    h = gpu_ctx.trace(g["rays"], bvh=1, split=True)
    check_hits(h, g["prim"], g["t"], g["u"], g["v"], exact=True)
    ha = gpu_ctx.trace(g["rays"], bvh=1, any_hit=True, split=True)
    assert np.array_equal(ha["prim"] != 0xFFFFFFFF, g["prim"] != 0xFFFFFFFF)


@pytest.mark.parametrize("mode", ["0", "3"])
@pytest.mark.parametrize("name", NAMES)
def test_wide_and_binary_walks_match_reference_golden(name, mode, bdpt, gpu_ctx, scenes, golden, monkeypatch):
    """bdpt_trace on BVH_FAST: the binary walk (default, BDPT_WIDE=0) and the optional 4-wide walk (BDPT_WIDE=3:
    closest- and any-hit queries through traverse_walk.cuh's run4) test the same candidates, so both must
    reproduce the reference."""
    g = golden(name)
    v, m, mats, cam = scenes[name]()
    monkeypatch.setenv("BDPT_WIDE", mode)
    gpu_ctx.upload(v, m, mats)
    monkeypatch.delenv("BDPT_WIDE")
    h = gpu_ctx.trace(g["rays"], bvh=bdpt.BVH_FAST)
    check_hits(h, g["prim"], g["t"], g["u"], g["v"], exact=True)
    ha = gpu_ctx.trace(g["rays"], bvh=bdpt.BVH_FAST, any_hit=True)
    assert np.array_equal(ha["prim"] != bdpt.MISS, g["prim"] != bdpt.MISS)
    gpu_ctx.stats_reset()
    gpu_ctx.trace(g["rays"], bvh=bdpt.BVH_FAST, stats=True)
    wide_nodes = gpu_ctx.stats()["node_visits"]
    assert wide_nodes > 0


def test_wide_walk_large_soup_vs_port(bdpt, gpu_ctx, monkeypatch):
    from oracle.oracle import Oracle
    from tools import testscenes as ts
    v, m, mats = ts.random_soup(60000, seed=21, extent=25.0, size=0.8)
    rays = ts.ray_batch(v, 200000, seed=77)
    o = Oracle("port").load(v, m, mats)
    prim, t, u, w = o.trace(rays, nthreads=8)
    visits = {}
    for mode in ("0", "3"):
        monkeypatch.setenv("BDPT_WIDE", mode)
        gpu_ctx.upload(v, m, mats)
        check_hits(gpu_ctx.trace(rays, bvh=bdpt.BVH_FAST), prim, t, u, w, exact=True)
        ha = gpu_ctx.trace(rays, bvh=bdpt.BVH_FAST, any_hit=True)
        assert np.array_equal(ha["prim"] != bdpt.MISS, prim != bdpt.MISS)
        gpu_ctx.stats_reset()
        gpu_ctx.trace(rays, bvh=bdpt.BVH_FAST, stats=True)
        visits[mode] = gpu_ctx.stats()["node_visits"]
    assert visits["3"] < 0.8 * visits["0"]                       # the 4-wide walk opens far fewer nodes


def test_host_buffer_batches_are_chunked_without_a_seam(bdpt, gpu_ctx):
    """bdpt_trace with host buffers runs 2^20-ray chunks on two streams (api.cu): a 2.3 M-ray batch (three chunks, the
    last one ragged) must equal the single-launch device-pointer path everywhere and the oracle across the seams."""
    import torch
    from tools import testscenes as ts
    v, m, mats = ts.random_soup(3000, seed=5, extent=12.0, size=0.9)
    n = (1 << 21) + 250001
    rays = ts.ray_batch(v, n, seed=99)
    gpu_ctx.upload(v, m, mats)
    for any_hit in (False, True):
        h = gpu_ctx.trace(rays, bvh=bdpt.BVH_FAST, any_hit=any_hit)
        d_rays = torch.from_numpy(np.ascontiguousarray(rays, dtype=np.float64)).cuda()
        d_hits = torch.zeros(n * bdpt.HIT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        gpu_ctx.trace_device(d_rays.data_ptr(), n, d_hits.data_ptr(), bvh=bdpt.BVH_FAST, any_hit=any_hit)
        hd = d_hits.cpu().numpy().view(bdpt.HIT_DTYPE)
        if any_hit:
            assert np.array_equal(h["prim"] != bdpt.MISS, hd["prim"] != bdpt.MISS)
        else:
            for f in ("prim", "t", "u", "v"):
                assert np.array_equal(h[f], hd[f]), f
    # pinned caller buffers (the asynchronous case) and a caller-provided output
    pr = torch.empty((n, 8), dtype=torch.float64, pin_memory=True)
    pr.numpy()[:] = rays
    po = torch.empty(n * bdpt.HIT_DTYPE.itemsize, dtype=torch.uint8, pin_memory=True).numpy().view(bdpt.HIT_DTYPE)
    hp = gpu_ctx.trace(pr.numpy(), bvh=bdpt.BVH_FAST, out=po)
    assert hp is po and np.array_equal(hp["prim"], gpu_ctx.trace(rays, bvh=bdpt.BVH_FAST)["prim"])
    assert gpu_ctx.last_device_ms() > 0
    o = Oracle("port").load(v, m, mats)
    for seam in (1 << 20, 1 << 21):
        sl = slice(seam - 500, seam + 500)
        prim, t, u, w = o.trace(rays[sl])
        check_hits(hp[sl], prim, t, u, w, exact=True)
    prim, t, u, w = o.trace(rays[-300:])
    check_hits(hp[-300:], prim, t, u, w, exact=True)
    with pytest.raises(ValueError):
        gpu_ctx.trace(rays[:10], bvh=bdpt.BVH_FAST, out=po)
