"""Parity at BASELINE.json's FULL sizes, where the CPU oracle would take an hour: size-independent properties.

config 2 (16,777,216 fixed-seed rays vs the glass-of-water BVH): the two tree flavours agree on every ray; an any-hit
query says "occluded" exactly where a closest-hit query finds something; re-tracing every hit ray with its interval
cut just behind / just in front of the hit finds the same triangle / a different or no triangle; a sample of the
batch is checked against the oracle itself.
config 3 (coffee_machine 1920x1080, depth 16): the film is additive over strata (camera.h:116-124 sums samples):
rendering strata {0,1} in one call equals rendering them in two calls and equals the sum of two sharded ranks.
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
    # a sample against the oracle itself (every 4099th ray: all three ray classes)
    sel = np.arange(0, n, 4099)
    o = Oracle("port").load(sf.verts, sf.mat_id, sf.material_list())
    prim, t, u, w = o.trace(rays[sel], nthreads=8)
    assert np.array_equal(fast["prim"][sel], prim) and np.array_equal(fast["t"][sel], t)


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
