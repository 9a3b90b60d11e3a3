"""Development aid: quality and speed of the three fast-tree builders on bench.py's scene (config 3) --
host binned SAH, device SAH over the Morton order, device LBVH (BDPT_GPU_TREE_LBVH=1)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
b = importlib.import_module("bidirectional-path-tracer_b200")
bench.ensure_scenes()
sf = b.SceneFile(os.environ.get("SCENE", bench.SCENE))
cam = sf.camera_with(bench.WIDTH, bench.HEIGHT, bench.SPP, bench.DEPTH)
K = int(sys.argv[1]) if len(sys.argv) > 1 else 4
for tag, gpu_tree, lbvh, sah in (("host_sah", False, "0", "1"), ("gpu_sah_all", True, "0", "1"), ("gpu_sah_bottoms", True, "0", "0"), ("gpu_lbvh", True, "1", "0")):
    os.environ["BDPT_GPU_TREE_LBVH"] = lbvh
    os.environ["BDPT_GPU_TREE_SAH"] = sah
    with b.Context(0) as ctx:
        ctx.upload(sf.verts, sf.mat_id, sf.materials, gpu_tree=gpu_tree)          # warm
        t0 = time.perf_counter(); ctx.upload(sf.verts, sf.mat_id, sf.materials, gpu_tree=gpu_tree); up = time.perf_counter() - t0
        ctx.film_clear(bench.WIDTH, bench.HEIGHT)
        ctx.stats_reset()
        ctx.render(cam, seed=1, bvh=b.BVH_FAST, strata=(0, 2, 1), clear=False, height=bench.HEIGHT, stats=True)
        s = ctx.stats()
        for k in range(2):
            ctx.render(cam, seed=1, bvh=b.BVH_FAST, strata=(4 * k + 4, 4 * k + 8, 1), clear=False, height=bench.HEIGHT)
        ms = 0.0
        for k in range(2, 2 + K):
            ctx.render(cam, seed=1, bvh=b.BVH_FAST, strata=(4 * k + 4, 4 * k + 8, 1), clear=False, height=bench.HEIGHT)
            ms += ctx.last_device_ms()
        print("%-15s upload %.1f ms  nodes %d depth %d | closest %.2f nodes %.2f tris /ray | shadow %.2f nodes %.2f tris /ray | %.1f Msamples/s %.2f ms/step"
              % (tag, up * 1e3, *ctx.bvh_info(1), s["nodes_closest"] / max(1, s["rays_closest"]), s["tris_closest"] / max(1, s["rays_closest"]),
                 s["nodes_shadow"] / max(1, s["rays_shadow"]), s["tris_shadow"] / max(1, s["rays_shadow"]),
                 bench.WIDTH * bench.HEIGHT * 4 * K / ms / 1e3, ms / K), flush=True)
