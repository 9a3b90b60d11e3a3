"""Development aid: does the walk's cost per node visit depend on whether the tree fits L1 / L2?  Same ray
generator on the Cornell room with spheres of growing tessellation."""
import importlib, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools import testscenes as ts
b = importlib.import_module("bidirectional-path-tracer_b200")
n = 4_000_000
with b.Context(0) as ctx:
    for steps in [(6, 8), (16, 32), (64, 128), (160, 320), (400, 800)]:
        v, m, mats, cam = ts.cornell(True, steps)
        rays = ts.ray_batch(v, n, seed=0x5EED0001, cam=cam)
        d_rays = None
        ctx.upload(v, m, mats)
        for any_hit in (False, True):
            for rep in range(3):
                ctx.trace(rays, bvh=1, any_hit=any_hit)
            ms = ctx.last_device_ms()
            ctx.stats_reset(); ctx.trace(rays, bvh=1, any_hit=any_hit, stats=True); s = ctx.stats()
            nv, tt = s["node_visits"] / n, s["tri_tests"] / n
            print("tris %8d any %d  Mrays/s %8.1f  nodes/ray %6.2f tris/ray %5.2f  Gnodes/s %6.2f  ns per (node+tri) per SM-lane-slot %.3f"
                  % (len(v), any_hit, n / ms / 1e3, nv, tt, nv * n / ms / 1e6, ms * 1e6 / (n * (nv + tt))), flush=True)
