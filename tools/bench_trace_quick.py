"""Quick device-timed trace microbench (development aid; bench.py is the contract)."""
import importlib, sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools import testscenes as ts
b = importlib.import_module("bidirectional-path-tracer_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
v, m, mats, cam = ts.cornell(True, (64, 128))
print("tris", len(v))
rays = ts.ray_batch(v, n, seed=0x5EED0001, cam=cam)
with b.Context(0) as ctx:
    t0 = time.time(); ctx.upload(v, m, mats); print("upload+build s", time.time() - t0, ctx.bvh_info(0), ctx.bvh_info(1))
    for bvh in (0, 1):
        for any_hit in (False, True):
            for rep in range(3):
                t0 = time.time(); h = ctx.trace(rays, bvh=bvh, any_hit=any_hit); wall = time.time() - t0
            ms = ctx.last_device_ms()
            print("bvh", bvh, "any", any_hit, "device ms %.3f  Mrays/s %.1f  e2e Mrays/s %.1f" % (ms, n / ms / 1e3, n / wall / 1e6))
        ctx.stats_reset(); ctx.trace(rays, bvh=bvh, stats=True); s = ctx.stats()
        print("  nodes/ray %.2f tris/ray %.2f" % (s["node_visits"] / n, s["tri_tests"] / n))
    h0 = ctx.trace(rays, bvh=0); h1 = ctx.trace(rays, bvh=1)
    print("fast==ref prim:", np.array_equal(h0["prim"], h1["prim"]), "t:", np.array_equal(h0["t"], h1["t"]))
