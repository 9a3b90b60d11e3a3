"""BDPT_WEIGHT_MIS (north_star item 4): balance-heuristic bidirectional path tracing with light-tracing
splats through film atomics.  Not a reference behaviour (the reference sums its strategies unweighted), so
the oracle is the port's restatement of the same estimator (oracle/port.cpp sample_mis), which tests/test_oracle.py
ties to the reference's own unbiased path tracer (camera::path_trace_color through oracle/_ref)."""
import numpy as np
import pytest

from oracle.oracle import Oracle, available, make_camera as ocam, RNG_PHILOX, WEIGHT_MIS

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    return float(np.abs(a - b).max() / max(1e-300, np.abs(b).max()))


@pytest.mark.parametrize("name", ["cornell_spheres", "cornell_box", "unit_room", "sphere_grid"])
@pytest.mark.parametrize("bvh", [0, 1])
def test_mis_film_equals_the_port_sample_for_sample(name, bvh, bdpt, gpu_ctx, scenes):
    v, m, mats, cam = scenes[name]()
    W, H, spp, depth = 40, 30, 4, 7
    ref = Oracle("port").load(v, m, mats).render(ocam(W, H, spp, depth, **cam), rng_mode=RNG_PHILOX, seed=42, nthreads=8, height=H,
                                                 weighting=WEIGHT_MIS)
    gpu_ctx.upload(v, m, mats)
    gpu_ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=42, bvh=bvh, height=H, weighting=bdpt.WEIGHT_MIS)
    film = gpu_ctx.film()
    assert ref.max() > 0
    assert rel_err(film, ref) < 1e-9          # fp64 atomics: only the order of the adds differs


def test_mis_film_is_independent_of_batching_and_sharding(bdpt, gpu_ctx, scenes):
    """Splats land in any pixel, from any rank: the per-pass light-path count in the weights is the frame's W * H no
    matter which strata or rows a call renders, so shards and row bands still add up to the whole frame."""
    v, m, mats, cam = scenes["cornell_spheres"]()
    gpu_ctx.upload(v, m, mats)
    c = bdpt.make_camera(32, 20, 16, 6, **cam)
    gpu_ctx.render(c, seed=3, height=20, bvh=1, weighting=bdpt.WEIGHT_MIS)
    whole = gpu_ctx.film().copy()
    gpu_ctx.film_clear(32, 20)
    for g in range(4):
        gpu_ctx.render(c, seed=3, strata=(g, 0, 4), batch_samples=777, clear=False, height=20, bvh=1, weighting=bdpt.WEIGHT_MIS)
    assert rel_err(gpu_ctx.film(), whole) < 1e-12
    gpu_ctx.film_clear(32, 20)
    gpu_ctx.render(c, seed=3, rows=(0, 7), clear=False, height=20, bvh=1, weighting=bdpt.WEIGHT_MIS)
    gpu_ctx.render(c, seed=3, rows=(7, 20), clear=False, height=20, bvh=1, weighting=bdpt.WEIGHT_MIS)
    assert rel_err(gpu_ctx.film(), whole) < 1e-12


def test_mis_converges_to_the_reference_path_tracer_with_less_error(bdpt, gpu_ctx, scenes):
    """Equal-spp relMSE against a 4,096-spp render: MIS must beat the reference's own unbiased estimator
    (path_trace_color, one light sample per bounce) -- and its mean must agree with it."""
    if not available("ref"):
        pytest.skip("oracle/_ref not shipped")
    v, m, mats, cam = scenes["cornell_box"]()
    W = H = 32
    depth, spp, seeds = 6, 16, 6
    gpu_ctx.upload(v, m, mats)
    gpu_ctx.render(bdpt.make_camera(W, H, 4096, depth, **cam), seed=99, bvh=1, height=H, weighting=bdpt.WEIGHT_MIS)
    truth = gpu_ctx.film() / 4096.0
    cpu = Oracle("ref").load(v, m, mats)

    def rel_mse(x):
        return float(np.mean((x - truth) ** 2 / (truth ** 2 + 1e-2)))

    e_mis, e_pt, mean_pt = [], [], []
    for k in range(seeds):
        gpu_ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=500 + k, bvh=1, height=H, weighting=bdpt.WEIGHT_MIS)
        e_mis.append(rel_mse(gpu_ctx.film() / spp))
        pt = cpu.render_pt(ocam(W, H, spp, depth, **cam), seed=700 + k, height=H) / spp
        e_pt.append(rel_mse(pt))
        mean_pt.append(pt.mean())
    print("relMSE at %d spp: MIS %.5f, reference path tracer %.5f" % (spp, np.mean(e_mis), np.mean(e_pt)))
    assert np.mean(e_mis) < np.mean(e_pt)
    assert abs(np.mean(mean_pt) / truth.mean() - 1.0) < 0.03
