"""Where the end-to-end time of a fresh context goes (host BVH builds, H2D, workspace allocation, render, D2H)."""
import importlib, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
pkg = importlib.import_module("bidirectional-path-tracer_b200")
bench.ensure_scenes()
sf = pkg.SceneFile(bench.SCENE)
cam = sf.camera_with(bench.WIDTH, bench.HEIGHT, bench.SPP, bench.DEPTH)
warm = pkg.Context(0); warm.upload(sf.verts, sf.mat_id, sf.materials); warm.render(cam, strata=(0, 1, 1), bvh=pkg.BVH_FAST, height=bench.HEIGHT); warm.close()
for rep in range(2):
    t = [time.perf_counter()]
    ctx = pkg.Context(0); t.append(time.perf_counter())
    ctx.upload(sf.verts, sf.mat_id, sf.materials); t.append(time.perf_counter())
    ctx.film_clear(bench.WIDTH, bench.HEIGHT); t.append(time.perf_counter())
    ctx.render(cam, strata=(0, 4, 1), bvh=pkg.BVH_FAST, height=bench.HEIGHT, clear=False); t.append(time.perf_counter())
    ctx.render(cam, strata=(4, 8, 1), bvh=pkg.BVH_FAST, height=bench.HEIGHT, clear=False); t.append(time.perf_counter())
    f = ctx.film(); t.append(time.perf_counter())
    ctx.close(); t.append(time.perf_counter())
    names = ["ctx_create", "upload(build+H2D)", "film_clear", "render#1(+workspace alloc)", "render#2", "film D2H", "close"]
    print({n: round(1e3 * (b - a), 1) for n, a, b in zip(names, t, t[1:])})
