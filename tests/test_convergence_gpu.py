"""Whole renders are Monte Carlo and cannot be bit-exact against the reference's mt19937 stream:
the gate is that the relative MSE against a high-spp render falls at the same RATE as the reference's own
output and is the same SIZE at equal spp.

  relMSE(x) = mean over pixels and channels of (x - ref)^2 / (ref^2 + 1e-2), on the linear averaged film,
              without the largest 2 % of the terms (fireflies of the unweighted estimator)
  ref       = 4096 spp (64 x 64 strata); the estimator is the same function in both arms and the GPU arm equals
              the oracle port sample for sample (tests/test_render_gpu.py), so the GPU renders it
  CPU arm   = the reference's own headers (oracle/_ref) when built, else the bit-identical port, mt19937 stream
  gate      = |slope_gpu - slope_cpu| <= 0.1 (log-log slope over spp 1, 4, 16, 64) and relMSE ratio at every spp
              within [0.8, 1.25] — SURVEY 8(d)'s numbers; each relMSE is the mean over SEEDS independent renders
"""
import numpy as np
import pytest

from oracle.oracle import Oracle, available, make_camera as ocam, RNG_MT19937

pytestmark = pytest.mark.gpu

W = H = 32
DEPTH = 6
SPPS = (1, 4, 16, 64)
SEEDS = 16


TRIM = 0.02


def rel_mse(film_sum, spp, ref):
    """The reference's estimator has the unweighted 1/dist^2 connection term (camera.h:473): its variance is
    unbounded, and a plain mean is ruled by one firefly.  As usual for relMSE the largest TRIM of the per-value
    errors are dropped — in both arms alike."""
    x = np.nan_to_num(film_sum / float(spp))
    e = np.sort(((x - ref) ** 2 / (ref ** 2 + 1e-2)).ravel())
    return float(np.mean(e[: int(len(e) * (1.0 - TRIM))]))


def test_relmse_converges_like_the_reference(bdpt, gpu_ctx, scenes):
    v, m, mats, cam = scenes["cornell_box"]()
    gpu_ctx.upload(v, m, mats)
    gpu_ctx.render(bdpt.make_camera(W, H, 4096, DEPTH, **cam), seed=987654321, bvh=bdpt.BVH_FAST, height=H)
    ref = np.nan_to_num(gpu_ctx.film() / 4096.0)
    assert ref.max() > 0
    cpu = Oracle("ref" if available("ref") else "port").load(v, m, mats)
    g, c = [], []
    for spp in SPPS:
        eg, ec = [], []
        for k in range(SEEDS):
            gpu_ctx.render(bdpt.make_camera(W, H, spp, DEPTH, **cam), seed=1000 + 17 * k + spp, bvh=bdpt.BVH_FAST, height=H)
            eg.append(rel_mse(gpu_ctx.film(), spp, ref))
            f = cpu.render(ocam(W, H, spp, DEPTH, **cam), rng_mode=RNG_MT19937, seed=5000 + 31 * k + spp, height=H)
            ec.append(rel_mse(f, spp, ref))
        g.append(np.mean(eg))
        c.append(np.mean(ec))
    lx = np.log(np.array(SPPS, dtype=np.float64))
    slope_g = np.polyfit(lx, np.log(g), 1)[0]
    slope_c = np.polyfit(lx, np.log(c), 1)[0]
    print("relMSE gpu", g, "cpu", c, "slopes", slope_g, slope_c)
    # both converge at all (not a clean power law at 1-4 spp: most one-sample pixels see no connection)
    assert g[-1] < 0.5 * g[1] and c[-1] < 0.5 * c[1]
    assert abs(slope_g - slope_c) <= 0.1
    for a, b in zip(g, c):
        assert 0.8 <= a / b <= 1.25, (g, c)


# ---------------------------------------------------------------------------------------------------------------
# The NAMED scenes against a CPU-rendered ground truth (north_star: "relative MSE against a high-spp reference CPU
# render must converge at the same rate as the reference's own output at equal spp").
#   ground truth = tests/golden/truth_<scene>.npz: 4,096 spp rendered by the reference's own headers (oracle/_ref),
#                  committed by tools/make_convergence_truth.py (reduced resolution: 4,096 spp on the CPU costs
#                  ~2 min per scene at 64x64 / 96x54 and hours at the configs' full resolutions)
#   GPU arm      = libbdpt, Philox stream; CPU arm = the port in mt19937 mode (bit-identical to oracle/_ref,
#                  tests/test_oracle.py), independent seeds, rendered side by side on the host cores
#   gates        = as above: |slope_gpu - slope_cpu| <= 0.1 over spp 1, 4, 16, 64; relMSE ratio in [0.8, 1.25] at every spp
import json
import os
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMED_SEEDS = 12


@pytest.fixture(scope="module")
def standins():
    from tools import make_standin_scenes as ms
    ms.generate(detail=2, quiet=True)


@pytest.mark.parametrize("name", ["glass_of_water", "coffee_machine"])
def test_named_scene_relmse_vs_cpu_ground_truth(name, bdpt, gpu_ctx, standins):
    from oracle.oracle import Camera as OCamera
    g = np.load(os.path.join(ROOT, "tests", "golden", "truth_%s.npz" % name))
    meta = json.loads(str(g["meta"]))
    ref = g["film"].astype(np.float64)
    Wn, Hn, depth = meta["width"], meta["height"], meta["depth"] or None
    assert ref.shape == (Hn, Wn, 3) and ref.max() > 0
    sf = bdpt.SceneFile(os.path.join(ROOT, meta["scene"]))
    gpu_ctx.upload(sf.verts, sf.mat_id, sf.materials)
    cpu = Oracle("port").load(sf.verts, sf.mat_id, sf.material_list())

    def cpu_render(args):
        spp, seed = args
        cam = OCamera.from_buffer_copy(bytes(sf.camera_with(Wn, Hn, spp, depth)))
        return rel_mse(cpu.render(cam, rng_mode=RNG_MT19937, seed=seed, height=Hn), int(np.sqrt(spp)) ** 2, ref)

    jobs = [(spp, 7000 + 131 * k + spp) for spp in SPPS for k in range(NAMED_SEEDS)]
    with ThreadPoolExecutor(max_workers=os.cpu_count() or 8) as pool:
        fut = pool.map(cpu_render, jobs)                       # the CPU arm runs while the GPU arm does
        gm = []
        for spp in SPPS:
            e = []
            for k in range(NAMED_SEEDS):
                gpu_ctx.render(sf.camera_with(Wn, Hn, spp, depth), seed=3000 + 17 * k + spp, bvh=bdpt.BVH_FAST, height=Hn)
                e.append(rel_mse(gpu_ctx.film(), spp, ref))
            gm.append(float(np.mean(e)))
        ce = list(fut)
    cpu.close()
    cm = [float(np.mean(ce[i * NAMED_SEEDS:(i + 1) * NAMED_SEEDS])) for i in range(len(SPPS))]
    lx = np.log(np.array(SPPS, dtype=np.float64))
    slope_g, slope_c = np.polyfit(lx, np.log(gm), 1)[0], np.polyfit(lx, np.log(cm), 1)[0]
    table = {"scene": meta["scene"], "resolution": [Wn, Hn], "depth": meta["depth"], "truth_spp": meta["spp"], "spp": list(SPPS),
             "relmse_gpu": gm, "relmse_cpu_reference": cm, "slope_gpu": slope_g, "slope_cpu": slope_c, "seeds": NAMED_SEEDS}
    print("relMSE", json.dumps(table))
    out_dir = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out_dir):
        json.dump(table, open(os.path.join(out_dir, "relmse_%s.json" % name), "w"))
    # both arms converge (slowly on glass-of-water: every surface but the lights is a delta material, so a pixel sees light
    # only through specular chains that happen to end on an emitter -- its relMSE is ruled by rare, bright samples)
    assert gm[-1] < 0.85 * gm[0] and cm[-1] < 0.85 * cm[0]
    assert abs(slope_g - slope_c) <= 0.1, table
    for a, b in zip(gm, cm):
        assert 0.8 <= a / b <= 1.25, table
