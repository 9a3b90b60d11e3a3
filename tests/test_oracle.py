"""The oracle is pinned here: the CPU restatement (oracle/port.cpp) against (a) the committed golden
vectors produced by the reference's own headers (tools/make_golden.py) and (b) the reference library
itself where it exists (oracle/_ref, build container and — as a prebuilt .so — the GPU box)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle.oracle import Oracle, available, make_camera

NAMES = ["cornell_spheres", "cornell_box", "unit_room"]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("name", NAMES)
def test_port_matches_golden_trace(name, scenes, golden):
    g = golden(name)
    v, m, mats, cam = scenes[name]()
    o = Oracle("port").load(v, m, mats)
    assert o.node_count() == int(g["nodes"])
    assert np.array_equal(o.leaf_order(), g["leaf_order"])
    prim, t, u, w = o.trace(g["rays"])
    assert np.array_equal(prim, g["prim"])           # bit-exact IDs
    assert np.array_equal(t, g["t"]) and np.array_equal(u, g["u"]) and np.array_equal(w, g["v"])
    assert np.array_equal(o.visible(g["seg_a"], g["seg_b"]), g["visible"])


@pytest.mark.parametrize("name", NAMES)
def test_port_matches_golden_render(name, scenes, golden):
    """mt19937 stream restarts per process in the reference; the port owns its engine, so in-process is fine."""
    g = golden(name)
    v, m, mats, cam = scenes[name]()
    W, H, spp, depth = (int(x) for x in g["render"])
    film = Oracle("port").load(v, m, mats).render(make_camera(W, H, spp, depth, **cam), height=H)
    assert np.array_equal(film, g["film"])           # bit-identical fp64 film


@pytest.mark.skipif(not available("ref"), reason="oracle/_ref not built")
def test_port_matches_reference_library_random_soup():
    from tools import testscenes as ts
    v, m, mats = ts.random_soup(3000, seed=3)
    rays = ts.ray_batch(v, 6000, seed=11)
    a = Oracle("port").load(v, m, mats)
    b = Oracle("ref").load(v, m, mats)
    assert np.array_equal(a.leaf_order(), b.leaf_order())
    for x, y in zip(a.trace(rays), b.trace(rays)):
        assert np.array_equal(x, y)


def test_colors_to_rgb8_known_values():
    film = np.array([[[0.0, 4.0, 1.0], [0.25 * 4, 1e9, -1.0]]])  # sums over 4 spp
    out = Oracle("port").colors_to_rgb8(film, 4)
    # avg 0 ->0 ; 1.0 -> clamp .999 -> sqrt -> 255 ; .25 -> .5*256=128 ; negative clamps to 0
    assert out.tolist() == [[[0, 255, 128], [128, 255, 0]]]


def test_philox_mode_statistically_equals_mt_mode(scenes):
    """The counter-based stream only changes the random numbers, not the estimator."""
    v, m, mats, cam = scenes["cornell_box"]()
    o = Oracle("port").load(v, m, mats)
    c = make_camera(12, 12, 64, 5, **cam)
    a = o.render(c, height=12).mean()
    from oracle.oracle import RNG_PHILOX
    b = o.render(c, rng_mode=RNG_PHILOX, seed=5, nthreads=4, height=12).mean()
    assert abs(a - b) / a < 0.06


def test_mis_estimator_agrees_with_the_reference_path_tracer():
    """the MIS-weighted estimator (port; the CUDA path equals it sample for sample, tests/test_mis_gpu.py)
    is unbiased: its image mean equals that of the reference's OWN path tracer, camera::path_trace_color
    (camera.h:255-292) run through oracle/_ref, within Monte-Carlo error -- while the reference's unweighted sum
    (camera.h:294-323), which counts a path once per strategy, is far off."""
    from oracle.oracle import WEIGHT_MIS, RNG_PHILOX, available, Oracle, make_camera
    if not available("ref"):
        pytest.skip("oracle/_ref not built")
    from tools import testscenes as ts
    for scene, tol in ((ts.cornell(False), 0.03), (ts.unit_room(seed=1, n_blobs=6), 0.05)):
        v, m, mats, cam = scene
        W = H = 16
        c = make_camera(W, H, 400, 6, **cam)
        mis = Oracle("port").load(v, m, mats).render(c, rng_mode=RNG_PHILOX, seed=3, nthreads=8, height=H, weighting=WEIGHT_MIS).mean()
        ref_sum = Oracle("port").load(v, m, mats).render(c, rng_mode=RNG_PHILOX, seed=3, nthreads=8, height=H).mean()
        pt = Oracle("ref").load(v, m, mats).render_pt(c, seed=5, height=H).mean()
        assert abs(mis / pt - 1.0) < tol, (mis, pt)
        assert abs(ref_sum / pt - 1.0) > 0.15, (ref_sum, pt)
