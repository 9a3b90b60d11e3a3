"""Diagnostic: per-ray traversal work extremes of one bench step (stats build of the kernels)."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
pkg = importlib.import_module("bidirectional-path-tracer_b200")
bench.ensure_scenes()
sf = pkg.SceneFile(bench.SCENE)
cam = sf.camera_with(bench.WIDTH, bench.HEIGHT, bench.SPP, bench.DEPTH)
with pkg.Context(0) as ctx:
    ctx.upload(sf.verts, sf.mat_id, sf.materials)
    for bvh in (pkg.BVH_FAST,):
        ctx.stats_reset()
        ctx.render(cam, seed=1, bvh=bvh, strata=(0, 1, 1), clear=True, height=bench.HEIGHT, stats=True)
        s = ctx.stats()
        print({k: s[k] for k in ("rays_closest", "rays_shadow", "nodes_closest", "tris_closest", "nodes_shadow", "tris_shadow",
                                 "max_work_closest", "max_work_shadow", "deferred_closest", "deferred_shadow")})
        ctx.stats_reset()
        ctx.render(cam, seed=1, bvh=bvh, strata=(1, 5, 1), clear=True, height=bench.HEIGHT, streams=1)
        print(ctx.kernel_times())
        print("bvh info fast", ctx.bvh_info(pkg.BVH_FAST), "ref", ctx.bvh_info(pkg.BVH_REFERENCE))
