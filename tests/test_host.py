"""CPU-side checks of the product library: it loads, exports every symbol include/bdpt.h declares,
fails loudly without a GPU, and its host BVH builder reproduces the reference topology."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "bdpt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bdpt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(bdpt):
    L = ctypes.CDLL(bdpt.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 18
    for s in syms:
        assert hasattr(L, s), s
    assert sorted(bdpt.EXPORTS) == syms


def test_struct_layouts_match_header(bdpt):
    assert ctypes.sizeof(bdpt.Material) == 40
    assert ctypes.sizeof(bdpt.Camera) == 144
    assert ctypes.sizeof(bdpt.RenderOpts) == 56
    assert ctypes.sizeof(bdpt.Stats) == 160   # 16 + the four remaining BvhStats counters
    assert bdpt.RAY_DTYPE.itemsize == 64 and bdpt.HIT_DTYPE.itemsize == 32


def test_no_cpu_fallback(bdpt):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(bdpt.BdptError, match="no CPU fallback"):
        bdpt.Context(0)


@pytest.mark.parametrize("name", ["cornell_spheres", "cornell_box", "unit_room"])
def test_host_bvh_matches_reference_topology(name, bdpt, scenes, golden):
    g = golden(name)
    v, m, mats, cam = scenes[name]()
    order, n_ref, n_fast = bdpt.host_bvh_build(v, fast=True)
    assert n_ref == int(g["nodes"])
    assert np.array_equal(order, g["leaf_order"])
    assert 1 <= n_fast <= len(v)


def test_host_bvh_duplicate_keys_and_tiny_scenes(bdpt):
    """std::sort's unstable order on equal keys must match the reference's (same libstdc++)."""
    from oracle.oracle import Oracle
    rng = np.random.default_rng(5)
    base = rng.uniform(-1, 1, (1, 9))
    for n in (1, 2, 3, 5, 64, 257):
        v = np.repeat(base, n, axis=0)
        v[:, 1::3] += np.arange(n)[:, None] % 3  # many equal x / z minima
        o = Oracle("port").load(v, np.zeros(n, np.uint32), [(0, (0.5, 0.5, 0.5), 0.0)])
        order, n_ref, _ = bdpt.host_bvh_build(v, fast=True)
        assert n_ref == o.node_count()
        assert np.array_equal(order, o.leaf_order())


def test_host_bvh_rejects_empty(bdpt):
    with pytest.raises(bdpt.BdptError):
        bdpt.host_bvh_build(np.zeros((0, 9)))


def test_host_bvh_large_tied_scene_matches_the_reference_build(bdpt):
    """Spans above 32768 triangles take the threaded sort (host_scene.cu, sort_like_std): its permutation must
    be std::sort's, ties included.  Quantised coordinates give thousands of equal keys per axis; checked against
    the reference's own bvh_node (oracle/_ref) and the single-threaded port."""
    from oracle.oracle import Oracle
    rng = np.random.default_rng(77)
    n = 100000
    c = np.round(rng.uniform(-8, 8, (n, 1, 3)) * 4) / 4                    # centres on a 0.25 grid: massive ties
    v = (c + np.round(rng.uniform(-0.5, 0.5, (n, 3, 3)) * 8) / 8).reshape(n, 9)
    mats = [(0, (0.5, 0.5, 0.5), 0.0)]
    order, n_ref, _ = bdpt.host_bvh_build(v, fast=False)
    for kind in ("ref", "port"):
        try:
            o = Oracle(kind).load(v, np.zeros(n, np.uint32), mats)
        except Exception:
            if kind == "ref":
                continue   # the reference build is absent on this box
            raise
        assert n_ref == o.node_count()
        assert np.array_equal(order, o.leaf_order()), kind


# ---------------------------------------------------------------- the fp32 pre-filter of the warp-synchronous walk
def _exact_mt(o, d, v0, e1, e2):
    """triangle.h:41-67 in fp64: (line hit?, t)."""
    pvec = np.cross(d, e2)
    det = float(np.dot(e1, pvec))
    if abs(det) < 1e-8:
        return False, 0.0
    inv = 1.0 / det
    tvec = o - v0
    u = float(np.dot(tvec, pvec)) * inv
    if u < 0.0 or u > 1.0:
        return False, 0.0
    qvec = np.cross(tvec, e1)
    v = float(np.dot(d, qvec)) * inv
    if v < 0.0 or u + v > 1.0:
        return False, 0.0
    return True, float(np.dot(e2, qvec)) * inv


def test_prefilter_is_conservative(bdpt):
    """csrc/traverse_w32.cuh: a triangle the fp32 filter REJECTS (0) is never one the reference's fp64 test accepts
    inside [tmin, best]; a triangle it calls CERTAIN (2) is always accepted, with tmin <= t <= t_hi (closest) or
    strictly inside the segment (any-hit).  Rays are aimed at, just inside and just outside edges and vertices, at
    grazing angles, at slivers, from near and far, across triangle sizes from 1e-4 to 10 of a 100-unit scene."""
    L = bdpt.lib()
    rng = np.random.default_rng(20261105)
    vmax = 60.0
    n_certain = n_reject = n_maybe = n_hits = 0
    hi = ctypes.c_float(0)
    offs = np.array([-1e-3, -1e-5, -1e-7, -1e-9, 0.0, 1e-9, 1e-7, 1e-5, 1e-3, 0.3, 0.5])
    for it in range(60000):
        s = 10.0 ** rng.uniform(-4, 1)
        v0 = rng.uniform(-50, 50, 3)
        e1 = rng.normal(size=3) * s
        e2 = rng.normal(size=3) * s
        if it % 7 == 0:                                    # sliver: e2 nearly parallel to e1
            e2 = e1 * rng.uniform(0.2, 2.0) + rng.normal(size=3) * s * 1e-5
        a, b = rng.choice(offs), rng.choice(offs)
        if it % 3 == 0:
            a = 1.0 - b + rng.choice(offs) * (1 if it % 2 else -1)      # around the u + v = 1 edge
        target = v0 + a * e1 + b * e2
        o = target + rng.normal(size=3) * 10.0 ** rng.uniform(-3, 2)
        if it % 11 == 0:                                   # grazing: origin almost in the triangle's plane
            n = np.cross(e1, e2)
            nn = np.linalg.norm(n)
            if nn > 0:
                o = o - (np.dot(o - v0, n / nn) * (1 - 10.0 ** rng.uniform(-8, -2))) * (n / nn)
        o = np.clip(o, -400, 400)
        d = target - o
        if it % 2:
            d = d / max(np.linalg.norm(d), 1e-300)
        d = d * 10.0 ** rng.uniform(-1, 1) if it % 5 == 0 else d
        o, d = o.astype(np.float32).astype(np.float64), d.astype(np.float32).astype(np.float64)   # as bdpt's fixed batches
        if np.any(d == 0):
            continue
        hit, t = _exact_mt(o, d, v0, e1, e2)
        tmin = 0.001
        tmax = np.inf if it % 4 else abs(t) * rng.choice([0.5, 1 - 1e-6, 1 - 1e-9, 1.0, 1 + 1e-9, 1 + 1e-6, 2.0]) + 1e-3
        best = tmax if it % 9 else min(tmax, abs(t) * rng.choice([0.999999, 1.0, 1.000001, 1.5]) + 1e-12)
        ray = np.array([*o, *d, tmin, tmax])
        tri = np.array([*v0, *e1, *e2])
        for any_hit in (0, 1):
            f = L.bdpt_test_filter32(ray.ctypes.data, float(best), tri.ctypes.data, vmax, any_hit, ctypes.byref(hi))
            accepted = hit and tmin <= t <= best
            if f == 0:
                n_reject += 1
                assert not accepted, (it, any_hit, o, d, v0, e1, e2, t, best)
            elif f == 2:
                n_certain += 1
                assert hit and t >= tmin, (it, any_hit, t)
                if any_hit:
                    assert tmin < t < tmax, (it, t, tmax)
                else:
                    assert t <= float(hi.value), (it, t, hi.value)
            else:
                n_maybe += 1
        n_hits += hit
    # the filter must also be SHARP, or the walk degenerates into fp64 tests: most decisions are not "maybe"
    assert n_certain > 1000 and n_reject > 10000, (n_certain, n_reject, n_maybe)
    print("prefilter: reject %d, maybe %d, certain %d, line hits %d" % (n_reject, n_maybe, n_certain, n_hits))
