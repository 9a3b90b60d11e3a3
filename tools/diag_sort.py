"""How much does ray ordering buy the traversal kernel?  Incoherent rays (uniform origins in the scene box,
uniform directions) against the coffee_machine stand-in (or, with the argument `grid50`, config 5's 10 M-triangle sphere
grid, whose trees do not fit the L2), traced as generated and after sorting by (direction octant, 30-bit Morton code of the origin)."""
import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
pkg = importlib.import_module("bidirectional-path-tracer_b200")
bench.ensure_scenes()
if len(sys.argv) > 1 and sys.argv[1].startswith("grid"):
    from tools import testscenes as ts
    class _S: pass
    sf = _S()
    sf.verts, sf.mat_id, sf.materials, _ = ts.sphere_grid(int(sys.argv[1][4:]))
    sf.verts = np.asarray(sf.verts)
else:
    sf = pkg.SceneFile(bench.SCENE)
n = 4_000_000
rng = np.random.default_rng(7)
lo, hi = sf.verts.reshape(-1, 3).min(0), sf.verts.reshape(-1, 3).max(0)
if len(sys.argv) > 1 and sys.argv[1].startswith("grid"): hi[1] = 3.0   # among the spheres, like the paths' bounce rays (the lights hang far above)
o = (lo + (hi - lo) * rng.random((n, 3))).astype(np.float32).astype(np.float64)
d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True); d = d.astype(np.float32).astype(np.float64)
rays = np.concatenate([o, d, np.full((n, 1), 0.001), np.full((n, 1), np.inf)], 1)

def part1by2(x):
    x = x.astype(np.uint64) & 0x3FF
    x = (x | (x << 16)) & 0x30000FF; x = (x | (x << 8)) & 0x300F00F; x = (x | (x << 4)) & 0x30C30C3; x = (x | (x << 2)) & 0x9249249
    return x
q = np.clip(((o - lo) / (hi - lo) * 1023).astype(np.int64), 0, 1023)
morton = part1by2(q[:, 0]) | (part1by2(q[:, 1]) << 1) | (part1by2(q[:, 2]) << 2)
octant = ((d[:, 0] < 0).astype(np.uint64) | ((d[:, 1] < 0).astype(np.uint64) << 1) | ((d[:, 2] < 0).astype(np.uint64) << 2))
orders = {"as generated": np.arange(n), "morton(origin)": np.argsort(morton, kind="stable"),
          "octant, morton": np.argsort((octant << 30) | morton, kind="stable"),
          "morton coarse 5 bits/axis": np.argsort(morton >> 15, kind="stable")}
with pkg.Context(0) as ctx:
    ctx.upload(sf.verts, sf.mat_id, sf.materials)
    for name, idx in orders.items():
        r = np.ascontiguousarray(rays[idx])
        best = 1e9
        for rep in range(3):
            ctx.trace(r, bvh=pkg.BVH_FAST); best = min(best, ctx.last_device_ms())
        print("%-28s %.3f ms  %.0f Mrays/s" % (name, best, n / best / 1e3))
