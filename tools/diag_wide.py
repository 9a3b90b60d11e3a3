"""Development aid: binary vs 4-wide walk (BDPT_WIDE=1) on the named scenes' stand-in geometry, device-timed."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools import testscenes as ts, make_standin_scenes as ms
b = importlib.import_module("bidirectional-path-tracer_b200")
ms.generate(detail=2, quiet=True)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
for name in ("coffee_machine", "glass-of-water"):
    sf = b.SceneFile(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes", name + ".yaml"))
    cam = dict(lookfrom=tuple(sf.camera.lookfrom), lookat=tuple(sf.camera.lookat), vfov=sf.camera.vfov)
    rays = ts.ray_batch(sf.verts, n, seed=0x5EED0001, cam=cam)
    with b.Context(0) as ctx:
        ctx.upload(sf.verts, sf.mat_id, sf.materials)
        href = ctx.trace(rays, bvh=0)
        best = min((ctx.trace(rays, bvh=0), ctx.last_device_ms())[1] for _ in range(3))
        print(name, "reference topology closest %.3f ms  %.0f Mrays/s" % (best, n / best / 1e3), flush=True)
        for any_hit in (False, True):
            best = 1e9
            for rep in range(4):
                h = ctx.trace(rays, bvh=1, any_hit=any_hit)
                best = min(best, ctx.last_device_ms())
            ctx.stats_reset(); ctx.trace(rays, bvh=1, any_hit=any_hit, stats=True); s = ctx.stats()
            ok = np.array_equal(h["prim"], href["prim"]) and np.array_equal(h["t"], href["t"]) if not any_hit else \
                np.array_equal(h["prim"] != b.MISS, ctx.trace(rays, bvh=0, any_hit=True)["prim"] != b.MISS)
            print(name, "wide" if os.environ.get("BDPT_WIDE") == "1" else "binary", "any" if any_hit else "closest",
                  "%.3f ms  %.0f Mrays/s  nodes/ray %.2f tris/ray %.2f  same answers: %s"
                  % (best, n / best / 1e3, s["node_visits"] / n, s["tri_tests"] / n, ok), flush=True)
