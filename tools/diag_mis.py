"""The two weightings side by side: device-timed Msamples/s of bench.py's step
(coffee_machine 1080p, 4 strata) under BDPT_WEIGHT_REF_SUM and BDPT_WEIGHT_MIS, and equal-spp relMSE on the Cornell
box against a 4,096-spp MIS render: MIS, the reference's unweighted sum (which converges to a different image: its
relMSE is dominated by that bias), and the reference's own unbiased path tracer on the CPU (oracle/_ref)."""
import importlib, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from tools import testscenes as ts
from oracle.oracle import Oracle, available, make_camera as ocam
b = importlib.import_module("bidirectional-path-tracer_b200")
out = {}
bench.ensure_scenes()
sf = b.SceneFile(bench.SCENE)
cam = sf.camera_with(bench.WIDTH, bench.HEIGHT, bench.SPP, bench.DEPTH)
with b.Context(0) as ctx:
    ctx.upload(sf.verts, sf.mat_id, sf.materials)
    for tag, w in (("ref_sum", b.WEIGHT_REF_SUM), ("mis", b.WEIGHT_MIS)):
        ctx.film_clear(bench.WIDTH, bench.HEIGHT)
        for s in range(2):
            ctx.render(cam, seed=1, bvh=b.BVH_FAST, strata=(4 * s, 4 * s + 4, 1), clear=False, height=bench.HEIGHT, weighting=w)
        ms, K = 0.0, 4
        ctx.stats_reset()
        for s in range(2, 2 + K):
            ctx.render(cam, seed=1, bvh=b.BVH_FAST, strata=(4 * s, 4 * s + 4, 1), clear=False, height=bench.HEIGHT, weighting=w)
            ms += ctx.last_device_ms()
        st = ctx.stats()
        out["coffee_machine_1080p_" + tag] = {"msamples_per_s": bench.WIDTH * bench.HEIGHT * 4 * K / ms / 1e3, "ms_per_step": ms / K,
                                              "shadow_rays_per_sample": st["rays_shadow"] / st["samples"],
                                              "closest_rays_per_sample": st["rays_closest"] / st["samples"]}
    v, m, mats, c = ts.cornell(False)
    W = H = 32
    depth, spp, seeds = 6, 16, 8
    ctx.upload(v, m, mats)
    ctx.render(b.make_camera(W, H, 4096, depth, **c), seed=99, bvh=1, height=H, weighting=b.WEIGHT_MIS)
    truth = ctx.film() / 4096.0
    rel = lambda x: float(np.mean((x - truth) ** 2 / (truth ** 2 + 1e-2)))
    e = {"mis": [], "ref_sum": [], "cpu_path_tracer": []}
    cpu = Oracle("ref").load(v, m, mats) if available("ref") else None
    for k in range(seeds):
        for tag, w in (("mis", b.WEIGHT_MIS), ("ref_sum", b.WEIGHT_REF_SUM)):
            ctx.render(b.make_camera(W, H, spp, depth, **c), seed=500 + k, bvh=1, height=H, weighting=w)
            e[tag].append(rel(ctx.film() / spp))
        if cpu:
            e["cpu_path_tracer"].append(rel(cpu.render_pt(ocam(W, H, spp, depth, **c), seed=700 + k, height=H) / spp))
    out["cornell_box_32x32_depth6_relmse_at_16spp_vs_4096spp_mis"] = {k: float(np.mean(x)) for k, x in e.items() if x}
print(json.dumps(out, indent=1))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/mis.json", "w"), indent=1)
