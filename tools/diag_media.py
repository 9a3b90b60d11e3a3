"""What the MEDIA / sphere flavours of the kernels cost at bench size: device-timed Msamples/s of bench.py's step (coffee_machine
1080p, 4 strata of every pixel) plain, with two constant_medium boxes (12 boundary triangles each), with three analytic spheres,
and with both -- under BDPT_WEIGHT_REF_SUM and BDPT_WEIGHT_MIS.  Writes gpurun_out/media.json."""
import importlib, json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from tools import testscenes as ts
b = importlib.import_module("bidirectional-path-tracer_b200")
bench.ensure_scenes()
sf = b.SceneFile(bench.SCENE)
cam = sf.camera_with(bench.WIDTH, bench.HEIGHT, bench.SPP, bench.DEPTH)
P = sf.verts.reshape(-1, 3)
lo, hi = P.min(0), P.max(0)
ext = hi - lo
c = (lo + hi) / 2
media = [(4.0 / ext.max(), (0.9, 0.9, 0.9), ts.box_mesh(tuple(c - 0.25 * ext), tuple(c + 0.25 * ext))),
         (1.5 / ext.max(), (0.2, 0.4, 0.9), ts.box_mesh(tuple(lo + 0.05 * ext), tuple(lo + ext * np.array([0.95, 0.35, 0.95]))))]
r = 0.08 * ext.max()
lamb = next(i for i, m in enumerate(sf.materials) if m.type == 0)
spheres = [(tuple(c + np.array([0.3, 0.2, 0.0]) * ext), r, lamb), (tuple(c + np.array([-0.3, 0.25, 0.1]) * ext), r, lamb),
           (tuple(c + np.array([0.0, 0.35, -0.3]) * ext), r, lamb)]
ball = [(4.0 / ext.max(), (0.9, 0.9, 0.9), ts.uv_sphere(tuple(c), 0.3 * ext.max(), 16, 32).reshape(-1, 9))]   # a large boundary: goes through its tree
out = {}
STREAMS = int(os.environ.get("STREAMS", "0"))        # 1: one worker, kernel classes' times add up to the step
ONLY = os.environ.get("ONLY", "")
with b.Context(0) as ctx:
    ctx.upload(sf.verts, sf.mat_id, sf.materials)
    for name, md, sp in (("plain", [], []), ("two_media", media, []), ("three_spheres", [], spheres), ("media_and_spheres", media, spheres),
                         ("ball_medium_%d_triangles" % len(ball[0][2]), ball, [])):
        if ONLY and not any(name.startswith(x) for x in ONLY.split(",")): continue
        ctx.set_media(md); ctx.set_spheres(sp)
        for tag, w in (("ref_sum", b.WEIGHT_REF_SUM), ("mis", b.WEIGHT_MIS)):
            ctx.film_clear(bench.WIDTH, bench.HEIGHT)
            for s in range(2):
                ctx.render(cam, seed=1, bvh=b.BVH_FAST, strata=(4 * s, 4 * s + 4, 1), clear=False, height=bench.HEIGHT, weighting=w, streams=STREAMS)
            ms, K = 0.0, 3
            ctx.stats_reset()
            for s in range(2, 2 + K):
                ctx.render(cam, seed=1, bvh=b.BVH_FAST, strata=(4 * s, 4 * s + 4, 1), clear=False, height=bench.HEIGHT, weighting=w, streams=STREAMS)
                ms += ctx.last_device_ms()
            st = ctx.stats()
            out[name + "_" + tag] = {"msamples_per_s": round(bench.WIDTH * bench.HEIGHT * 4 * K / ms / 1e3, 1), "ms_per_step": round(ms / K, 2),
                                     "closest_rays_per_sample": round(st["rays_closest"] / st["samples"], 3),
                                     "shadow_rays_per_sample": round(st["rays_shadow"] / st["samples"], 3),
                                     "film_mean": float(np.nan_to_num(ctx.film()).mean())}
            out[name + "_" + tag]["kernel_ms_all_calls"] = {k: round(v[0], 1) for k, v in ctx.kernel_times().items() if v[0] > 0}
            print(name, tag, out[name + "_" + tag], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/media.json", "w"), indent=1)
