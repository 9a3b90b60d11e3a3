"""Development aid: what the scene image buys at config 5's scale (10 M triangles): cold upload vs image upload."""
import importlib, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools import testscenes as ts
b = importlib.import_module("bidirectional-path-tracer_b200")
g = int(sys.argv[1]) if len(sys.argv) > 1 else 50
v, m, mats, camkw = ts.sphere_grid(g)
print("triangles", len(v), flush=True)
path = "/tmp/grid.bdptimg"
with b.Context(0) as ctx:
    t0 = time.perf_counter(); ctx.upload(v, m, mats); t_cold = time.perf_counter() - t0
    rays = ts.ray_batch(v, 200000, seed=3)
    want = ctx.trace(rays, bvh=b.BVH_FAST)
    t0 = time.perf_counter(); img = b.SceneImage.build(v, m, mats); t_build = time.perf_counter() - t0
    t0 = time.perf_counter(); img.save(path); t_save = time.perf_counter() - t0
    nbytes = img.nbytes
    t0 = time.perf_counter(); ctx.upload_image(img); t_mem = time.perf_counter() - t0
    img.close()
    t0 = time.perf_counter(); mp = b.SceneImage.map(path); t_map = time.perf_counter() - t0
    t0 = time.perf_counter(); mp.info(verify=True); t_verify = time.perf_counter() - t0
    t0 = time.perf_counter(); ctx.upload_image(mp); t_file = time.perf_counter() - t0
    got = ctx.trace(rays, bvh=b.BVH_FAST)
    same = all(np.array_equal(want[f], got[f]) for f in ("prim", "t", "u", "v"))
    mp.close()
    os.remove(path)
    print({"triangles": len(v), "image_MB": nbytes / 1e6, "cold_upload_s": t_cold, "image_build_s": t_build, "save_s": t_save,
           "upload_from_memory_s": t_mem, "map_s": t_map, "verify_hash_s": t_verify, "upload_from_mapped_file_s": t_file,
           "same_hits_after_image_upload": same})
