import importlib, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tools import testscenes as ts
pkg = importlib.import_module("bidirectional-path-tracer_b200")
v, m, mats, cam = ts.cornell(True, (8, 12))
c = pkg.make_camera(24, 24, 9, 10, **cam)
with pkg.Context(0) as ctx:
    ctx.upload(v, m, mats)
    os.environ.pop("BDPT_WORK_LIMIT", None); os.environ.pop("BDPT_WORK_LIMIT_SHADOW", None)
    ctx.render(c, seed=11, bvh=pkg.BVH_FAST, height=24)
    whole = ctx.film()
    for lt, ls in (("8", "0"), ("0", "8"), ("40", "0"), ("0", "40")):
        os.environ["BDPT_WORK_LIMIT"] = lt; os.environ["BDPT_WORK_LIMIT_SHADOW"] = ls
        ctx.stats_reset()
        ctx.render(c, seed=11, bvh=pkg.BVH_FAST, height=24)
        f = ctx.film()
        d = np.abs(f - whole).max(axis=2)
        st = ctx.stats()
        print("trace limit", lt, "shadow limit", ls, "max diff", d.max(), "pixels differing", int((d > 1e-12).sum()),
              "sum ratio", f.sum() / whole.sum(), "deferred", st["deferred_closest"], st["deferred_shadow"], "rays", st["rays_closest"], st["rays_shadow"])
