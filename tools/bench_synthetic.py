#!/usr/bin/env python
"""BASELINE.json config 5: synthetic tessellated scene (~10 M triangles, mixed dielectric / fuzzy metal spheres,
depth 32, 1080p, 64 spp) — the divergence / HBM stress case: its BVH (~0.4 GB of nodes + 0.8 GB of triangles)
does not fit the 126 MB L2.  Prints one JSON line (not the bench.py contract; recorded in BASELINE.md).

  python tools/bench_synthetic.py [grid=50] [steps=4] [gpu_tree=0]      # grid x grid spheres of 4,000 triangles
"""
import importlib, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools import testscenes as ts

g = int(sys.argv[1]) if len(sys.argv) > 1 else 50
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
gpu_tree = bool(int(sys.argv[3])) if len(sys.argv) > 3 else None
W, H, SPP, DEPTH, STRATA = 1920, 1080, 64, 32, 2
pkg = importlib.import_module("bidirectional-path-tracer_b200")
t0 = time.perf_counter()
v, m, mats, camkw = ts.sphere_grid(g)
t_gen = time.perf_counter() - t0
cam = pkg.make_camera(W, H, SPP, DEPTH, **camkw)
with pkg.Context(0) as ctx:
    t0 = time.perf_counter()
    ctx.upload(v, m, mats, gpu_tree=gpu_tree)
    t_up = time.perf_counter() - t0
    ctx.film_clear(W, H)
    for s in range(2):   # warm-up
        ctx.render(cam, seed=1, bvh=pkg.BVH_FAST, strata=(s * STRATA, (s + 1) * STRATA, 1), clear=False, height=H)
    ctx.stats_reset()
    ms = 0.0
    for s in range(steps):
        ctx.render(cam, seed=1, bvh=pkg.BVH_FAST, strata=((2 + s) * STRATA, (3 + s) * STRATA, 1), clear=False, height=H)
        ms += ctx.last_device_ms()
    st = ctx.stats()
    ctx.stats_reset()
    ctx.render(cam, seed=1, bvh=pkg.BVH_FAST, strata=(0, 1, 1), clear=True, height=H, stats=True)
    s2 = ctx.stats()
    samples = W * H * STRATA * steps
    line = {"workload": "synthetic sphere grid %dx%d, %d triangles, %dx%d, %d spp, depth %d; step = %d strata" % (g, g, len(v), W, H, SPP, DEPTH, STRATA),
            "msamples_per_s": samples / ms / 1e3, "mrays_per_s": (st["rays_closest"] + st["rays_shadow"]) / ms / 1e3,
            "ms_per_step": ms / steps, "rays_per_sample": {"closest": st["rays_closest"] / st["samples"], "shadow": st["rays_shadow"] / st["samples"]},
            "nodes_per_closest_ray": s2["nodes_closest"] / max(1, s2["rays_closest"]), "tris_per_closest_ray": s2["tris_closest"] / max(1, s2["rays_closest"]),
            "bvh_nodes": {"fast": ctx.bvh_info(pkg.BVH_FAST), "reference": ctx.bvh_info(pkg.BVH_REFERENCE)},
            "fast_tree": "host, binned SAH" if gpu_tree is False else "device, binned SAH", "host_s": {"generate": t_gen, "upload_and_both_bvh_builds": t_up}, "deferred_closest": st["deferred_closest"]}
    print(json.dumps(line))
