"""Parity at BASELINE.json's FULL sizes.

config 2 (16,777,216 fixed-seed rays vs the glass-of-water BVH): EVERY ray is compared with the CPU oracle (the port,
pinned bit-identical to the reference's own headers; ~0.65 Mrays/s on 8 threads, so the whole batch costs well under a
minute of host time): prim IDs equal on 100 % of the rays and t, u, v bit-equal, for both tree flavours.  On top of
that, size-independent properties: an any-hit query says "occluded" exactly where a closest-hit query finds something;
re-tracing every hit ray with its interval cut just behind / just in front of the hit finds the same triangle / nothing.
config 3 (coffee_machine 1920x1080, depth 16): the film is additive over strata (camera.h:116-124 sums samples):
rendering strata {0,1} in one call equals rendering them in two calls and equals the sum of two sharded ranks.
config 5 (10,000,008 triangles): prim IDs, t, u, v of a 2^20-ray sub-sample (camera / incoherent / shadow-segment thirds)
against the oracle's own tree over the same 10 M triangles, both fast-tree builders.
"""
import os

import numpy as np
import pytest

from oracle.oracle import Oracle

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def standins():
    from tools import make_standin_scenes as ms
    ms.generate(detail=2, quiet=True)


def test_config2_full_batch_properties(bdpt, gpu_ctx, standins):
    from tools import testscenes as ts
    sf = bdpt.SceneFile(os.path.join(ROOT, "scenes", "glass-of-water.yaml"))
    cam = {"lookfrom": tuple(sf.camera.lookfrom), "lookat": tuple(sf.camera.lookat), "vfov": sf.camera.vfov}
    n = 1 << 24
    rays = ts.ray_batch(sf.verts, n, seed=0x5EED0001, cam=cam, width=512, height=512)
    assert len(rays) == n
    gpu_ctx.upload(sf.verts, sf.mat_id, sf.materials)
    fast = gpu_ctx.trace(rays, bvh=bdpt.BVH_FAST)
    ref = gpu_ctx.trace(rays, bvh=bdpt.BVH_REFERENCE)
    for f in ("prim", "t", "u", "v"):
        assert np.array_equal(fast[f], ref[f]), f                     # bit-equal between the flavours, all 16.8 M rays
    hit = fast["prim"] != bdpt.MISS
    assert 0.3 < hit.mean() < 0.7
    anyh = gpu_ctx.trace(rays, bvh=bdpt.BVH_FAST, any_hit=True)
    assert np.array_equal(anyh["prim"] != bdpt.MISS, hit)
    # interval cut just behind the hit: same triangle, same t; just in front of it: that hit is gone
    idx = np.flatnonzero(hit)[:: 8]
    behind = rays[idx].copy()
    behind[:, 7] = fast["t"][idx] * (1 + 1e-6) + 1e-9
    hb = gpu_ctx.trace(behind, bvh=bdpt.BVH_FAST)
    assert np.array_equal(hb["prim"], fast["prim"][idx]) and np.array_equal(hb["t"], fast["t"][idx])
    front = rays[idx].copy()
    front[:, 7] = fast["t"][idx] * (1 - 1e-6) - 1e-9
    ok = front[:, 7] > front[:, 6]
    hf = gpu_ctx.trace(front[ok], bvh=bdpt.BVH_FAST)
    assert np.all(hf["prim"] == bdpt.MISS)                            # nothing lies in front of a CLOSEST hit
    # the whole batch against the oracle itself: BASELINE.json "bit-exact prim IDs" on all 16,777,216 rays
    o = Oracle("port").load(sf.verts, sf.mat_id, sf.material_list())
    prim, t, u, w = o.trace(rays, nthreads=os.cpu_count() or 8)
    o.close()
    bad = np.flatnonzero(fast["prim"] != prim)
    assert len(bad) == 0, "prim IDs differ from the oracle on %d of %d rays, first %r" % (len(bad), n, bad[:8])
    assert np.array_equal(fast["t"], t) and np.array_equal(fast["u"], u) and np.array_equal(fast["v"], w)
    assert np.array_equal(ref["prim"], prim) and np.array_equal(ref["t"], t)


def test_config3_film_is_additive_over_strata_at_full_size(bdpt, gpu_ctx, standins):
    sf = bdpt.SceneFile(os.path.join(ROOT, "scenes", "coffee_machine.yaml"))
    W, H = 1920, 1080
    cam = sf.camera_with(W, H, 256, 16)
    gpu_ctx.upload(sf.verts, sf.mat_id, sf.materials)

    def film(strata_list):
        gpu_ctx.film_clear(W, H)
        for s in strata_list:
            gpu_ctx.render(cam, seed=5, bvh=bdpt.BVH_FAST, strata=s, clear=False, height=H)
        return gpu_ctx.film()

    whole = film([(0, 2, 1)])
    assert np.isfinite(whole).all() and whole.sum() > 0
    scale = np.abs(whole).max()
    two_calls = film([(0, 1, 1), (1, 2, 1)])
    assert np.abs(two_calls - whole).max() / scale < 1e-12            # fp64 atomics: summation order only
    sharded = film([(0, 2, 2), (1, 2, 2)])                            # what rank 0 and rank 1 of 2 would render (begin, end, step)
    assert np.abs(sharded - whole).max() / scale < 1e-12
    st = gpu_ctx.stats()
    assert st["samples"] > 0


def test_config5_ten_million_triangles_subsample_vs_oracle(bdpt, gpu_ctx):
    """at 10 M triangles the trees are out of L2 and the fp32 boxes are ~1e-5 of the extent wide, which is
    where a conservative-culling mistake would show.  2^20 rays against the oracle's own median-split tree."""
    from tools import testscenes as ts
    v, m, mats, camkw = ts.sphere_grid(50)
    assert len(v) == 10_000_008
    rays = ts.ray_batch(v, 1 << 20, seed=0xB200, cam=camkw, width=1920, height=1080)
    o = Oracle("port").load(v, m, mats)
    prim, t, u, w = o.trace(rays, nthreads=os.cpu_count() or 8)
    o.close()
    assert 0.3 < (prim != bdpt.MISS).mean() < 0.9
    for gpu_tree in (False, True):
        gpu_ctx.upload(v, m, mats, gpu_tree=gpu_tree)
        for bvh in ((bdpt.BVH_FAST, bdpt.BVH_REFERENCE) if not gpu_tree else (bdpt.BVH_FAST,)):
            h = gpu_ctx.trace(rays, bvh=bvh)
            bad = np.flatnonzero(h["prim"] != prim)
            assert len(bad) == 0, "gpu_tree=%r bvh=%d: %d rays differ, first %r" % (gpu_tree, bvh, len(bad), bad[:8])
            assert np.array_equal(h["t"], t) and np.array_equal(h["u"], u) and np.array_equal(h["v"], w)
    # any-hit agrees with closest-hit about occlusion on the shadow-segment third
    seg = rays[-(1 << 18):]
    a = gpu_ctx.trace(seg, bvh=bdpt.BVH_FAST, any_hit=True)
    assert np.array_equal(a["prim"] != bdpt.MISS, prim[-(1 << 18):] != bdpt.MISS)
