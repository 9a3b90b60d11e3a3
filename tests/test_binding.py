"""The drop-in proof: tests/binding/bdpt_binding.h -- the one file a maintainer
adds to the reference -- compiled together with the reference's REAL headers and linked against libbdpt.so
(oracle/Makefile: _ref/binding_demo, built where /root/reference exists; the GPU box runs the prebuilt binary).
The binary rebuilds a scene with the reference's own triangle / material / camera classes and renders it through
the binding; the PNG must equal, byte for byte, the one the ctypes binding writes for the same soup, camera and seed
(rgb8 is integer output; the film differs at most in fp64 summation order, ~1e-16 relative)."""
import os
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEMO = os.path.join(ROOT, "oracle", "_ref", "binding_demo")
needs_demo = pytest.mark.skipif(not os.path.exists(DEMO), reason="oracle/_ref/binding_demo not built (needs /root/reference at build time)")


def write_scene_bin(path, v, m, mats, aspect, W, spp, depth, cam):
    v = np.ascontiguousarray(v, np.float64).reshape(-1, 9)
    m = np.ascontiguousarray(m, np.uint32)
    with open(path, "wb") as f:
        f.write(struct.pack("<QII", len(v), len(mats), 0))
        f.write(v.tobytes())
        f.write(m.tobytes())
        for t, rgb, p in mats:
            f.write(struct.pack("<II4d", int(t), 0, float(rgb[0]), float(rgb[1]), float(rgb[2]), float(p)))
        bg = cam.get("background", (0.0, 0.0, 0.0))
        vup = cam.get("vup", (0.0, 1.0, 0.0))
        f.write(struct.pack("<d4i", aspect, W, spp, depth, 0))
        f.write(struct.pack("<3d", *bg))
        f.write(struct.pack("<d", cam.get("vfov", 90.0)))
        f.write(struct.pack("<3d", *cam["lookfrom"]))
        f.write(struct.pack("<3d", *cam["lookat"]))
        f.write(struct.pack("<3d", *vup))


@needs_demo
def test_binding_demo_is_linked_against_the_product_and_fails_cleanly_without_a_scene(tmp_path):
    """CPU part: the TU compiled against the reference's headers, resolves libbdpt.so through its rpath, and reports
    errors the reference's way (message + exit 1, main.cpp:72-75)."""
    ldd = subprocess.run(["ldd", DEMO], capture_output=True, text=True).stdout
    assert "libbdpt.so" in ldd and "not found" not in ldd
    r = subprocess.run([DEMO, str(tmp_path / "missing.bin"), str(tmp_path / "o.png")], capture_output=True, text=True)
    assert r.returncode == 1 and "cannot open scene.bin" in r.stderr


@needs_demo
@pytest.mark.gpu
@pytest.mark.parametrize("name,W,H,spp,depth", [("cornell_spheres", 48, 32, 9, 8), ("unit_room", 40, 40, 5, 6)])
def test_reference_side_binding_renders_the_same_png_as_the_ctypes_binding(name, W, H, spp, depth, bdpt, scenes, tmp_path):
    v, m, mats, cam = scenes[name]()
    write_scene_bin(tmp_path / "scene.bin", v, m, mats, bdpt.aspect_for(W, H), W, spp, depth, cam)
    out = tmp_path / "binding.png"
    r = subprocess.run([DEMO, str(tmp_path / "scene.bin"), str(out), "77"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "%d triangles" % len(np.asarray(v).reshape(-1, 9)) in r.stdout
    with bdpt.Context(0) as ctx:
        ctx.upload(v, m, mats)
        ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=77, bvh=bdpt.BVH_FAST, height=H)
        n = int(np.sqrt(spp))
        rgb = ctx.film_rgb8(n * n)                      # camera.h:139-140: the average is over sqrt_spp^2
    assert rgb.any()
    ref_png = tmp_path / "ctypes.png"
    bdpt.write_png(str(ref_png), rgb)
    assert out.read_bytes() == ref_png.read_bytes()
