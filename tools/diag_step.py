"""Development aid: device-timed Msamples/s of bench.py's step (config 3, scene resident) under the current
environment (BDPT_WALK_SLICE, BDPT_MIN_IDLE, ... are read by bdpt_render).  bench.py is the contract."""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
b = importlib.import_module("bidirectional-path-tracer_b200")
bench.ensure_scenes()
sf = b.SceneFile(bench.SCENE)
cam = sf.camera_with(bench.WIDTH, bench.HEIGHT, bench.SPP, bench.DEPTH)
K = int(sys.argv[1]) if len(sys.argv) > 1 else 6
STREAMS, BATCH = int(os.environ.get("STREAMS", "0")), int(os.environ.get("BATCH", "0"))
with b.Context(0) as ctx:
    if os.environ.get("USE_IMAGE"):
        with b.SceneImage.build(sf.verts, sf.mat_id, sf.materials) as img:
            ctx.upload_image(img)
    else:
        ctx.upload(sf.verts, sf.mat_id, sf.materials)
    ctx.film_clear(bench.WIDTH, bench.HEIGHT)
    for s in range(3):
        ctx.render(cam, seed=1, bvh=b.BVH_FAST, strata=(4 * s, 4 * s + 4, 1), clear=False, height=bench.HEIGHT, streams=STREAMS, batch_samples=BATCH)
    ms = 0.0
    for s in range(3, 3 + K):
        ctx.render(cam, seed=1, bvh=b.BVH_FAST, strata=(4 * s, 4 * s + 4, 1), clear=False, height=bench.HEIGHT, streams=STREAMS, batch_samples=BATCH)
        ms += ctx.last_device_ms()
    kt = ctx.kernel_times()
    print(os.environ.get("TAG", ""), "%.1f Msamples/s  %.2f ms/step" % (bench.WIDTH * bench.HEIGHT * 4 * K / ms / 1e3, ms / K),
          {k: round(v[0], 1) for k, v in kt.items() if v[0] > 0}, flush=True)
