"""Textures (/root/reference/src/materials/textures/texture.h:28-87, perlin.h): checker, image and Perlin-noise
albedo / emission.  The reference's loader can build none of them, so the pin is the reference's own texture CLASSES driven
through oracle/_ref (tools/make_golden.py --textured -> tests/golden/cornell_textured.npz): texture::value at 4,000 points,
the image bytes as rtw_image holds them after its float round trip, a noise texture's perlin tables, and a whole render."""
import numpy as np
import pytest

from oracle import oracle as O
from tools import testscenes as ts


def textured_scene(golden, noise=True):
    g = golden("cornell_textured")
    v, m, mats, cam, texs = ts.cornell_textured(noise=noise)
    texs[1] = dict(texs[1], data=g["image_bytes"])             # what the reference's loader makes of the raw bytes
    if noise:
        texs[3] = dict(texs[3], data=g["noise_tables"])        # the tables the reference's perlin drew
    return g, v, m, mats, cam, texs


def test_port_texture_values_equal_the_reference_classes(golden):
    g, v, m, mats, cam, texs = textured_scene(golden)
    tarr, blob = O.make_textures(texs)
    port = O.Oracle("port").load(v, m, mats, textures=(tarr, len(texs), blob))
    for k in range(len(texs)):
        assert np.array_equal(port.texture_value(k, g["uvp"]), g["values"][k]), k      # bit-equal: same fp64 operations


def test_port_textured_render_equals_the_reference(golden):
    g, v, m, mats, cam, texs = textured_scene(golden, noise=False)
    tarr, blob = O.make_textures(texs)
    W, H, spp, depth = (int(x) for x in g["render"])
    film = O.Oracle("port").load(v, m, mats, textures=(tarr, len(texs), blob)).render(O.make_camera(W, H, spp, depth, **cam), height=H)
    assert np.array_equal(film, g["film"])


@pytest.mark.gpu
@pytest.mark.parametrize("weighting", [0, 1])
@pytest.mark.parametrize("bvh", [0, 1])
def test_gpu_textured_film_equals_the_port(bvh, weighting, bdpt, gpu_ctx, golden):
    g, v, m, mats, cam, texs = textured_scene(golden)
    tarr, blob = O.make_textures(texs)
    W, H, spp, depth = 36, 28, 4, 7
    ref = O.Oracle("port").load(v, m, mats, textures=(tarr, len(texs), blob)).render(
        O.make_camera(W, H, spp, depth, **cam), rng_mode=O.RNG_PHILOX, seed=11, nthreads=8, height=H, weighting=weighting)
    try:
        gpu_ctx.upload(v, m, mats)
        gpu_ctx.set_textures(texs)
        gpu_ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=11, bvh=bvh, height=H, weighting=weighting)
        film = gpu_ctx.film()
        assert ref.max() > 0
        # the noise texture's sine is libm's on the host and CUDA's on the device (<= 2 ulp apart): 1e-9 of the peak holds both
        assert np.abs(film - ref).max() / np.abs(ref).max() < 1e-9
        plain = [(t, rgb, p) for t, rgb, p, _ in mats]
        gpu_ctx.upload(v, m, plain)
        gpu_ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=11, bvh=bvh, height=H, weighting=weighting)
        assert np.abs(gpu_ctx.film() - ref).max() / np.abs(ref).max() > 1e-3      # the textures really were applied
    finally:
        gpu_ctx.set_textures([])


@pytest.mark.gpu
@pytest.mark.parametrize("weighting", [0, 1])
def test_gpu_film_with_textures_media_and_spheres_together_equals_the_port(weighting, bdpt, gpu_ctx, golden):
    """Every extension of the scene surface at once (the TEX + MEDIA kernel flavours under either weighting): textured room, two media,
    two analytic spheres of the scene's own materials."""
    g, v, m, mats, cam, texs = textured_scene(golden)
    tarr, blob = O.make_textures(texs)
    media = ts.cornell_media()
    lamb = next(i for i, mt in enumerate(mats) if mt[0] == 0)
    spheres = [((190.0, 90.0, 190.0), 90.0, lamb), ((400.0, 300.0, 350.0), 70.0, 3)]   # the second one wears a texture (u = v = 0 on a sphere prim)
    W, H, spp, depth = 36, 28, 4, 7
    port = O.Oracle("port").load(v, m, mats, textures=(tarr, len(texs), blob)).set_media(media).set_spheres(spheres)
    ref = port.render(O.make_camera(W, H, spp, depth, **cam), rng_mode=O.RNG_PHILOX, seed=17, nthreads=8, height=H, weighting=weighting)
    try:
        gpu_ctx.upload(v, m, mats)
        gpu_ctx.set_textures(texs); gpu_ctx.set_media(media); gpu_ctx.set_spheres(spheres)
        gpu_ctx.render(bdpt.make_camera(W, H, spp, depth, **cam), seed=17, bvh=1, height=H, weighting=weighting)
        assert ref.max() > 0
        assert np.abs(gpu_ctx.film() - ref).max() / np.abs(ref).max() < 1e-9
    finally:
        gpu_ctx.set_textures([]); gpu_ctx.set_media([]); gpu_ctx.set_spheres([])


@pytest.mark.gpu
def test_set_textures_rejects_bad_tables(bdpt, gpu_ctx):
    with pytest.raises(bdpt.BdptError):
        gpu_ctx.set_textures([dict(type=bdpt.TEX_IMAGE, width=100, height=100, data=np.zeros(12, np.uint8))])
    with pytest.raises(bdpt.BdptError):
        gpu_ctx.set_textures([dict(type=bdpt.TEX_CHECKER, scale=0.0)])
    with pytest.raises(bdpt.BdptError):
        gpu_ctx.set_textures([dict(type=7)])


def test_yaml_texture_extension_builds_the_reference_objects(bdpt, golden, tmp_path):
    """host/scene_loader.cpp: a `texture:` map in a lambertian / light material.  The noise texture's perlin tables are drawn
    the way perlin.h:6-14,68-75 draws them from the head of the reference's mt19937 stream, the image goes through
    rtw_image's float round trip: both must equal what the reference's own objects held (golden, minted through oracle/_ref)."""
    g = golden("cornell_textured")
    _, _, _, _, texs = ts.cornell_textured(noise=True)
    img = texs[1]["data"]
    with open(tmp_path / "wall.ppm", "wb") as f:
        f.write(b"P6\n12 9\n255\n" + img.tobytes())
    (tmp_path / "scene.yaml").write_text("""
camera:
  resolution: [32, 32]
materials:
  floor: {type: lambertian, texture: {type: checker, scale: 60.0, even: [0.8, 0.8, 0.8], odd: [0.15, 0.2, 0.6]}}
  wall: {type: lambertian, texture: {type: image, file: wall.ppm}}
  lamp: {type: light, texture: {type: checker, scale: 25.0, even: [15, 15, 15], odd: [15, 6, 2]}}
  marble: {type: lambertian, texture: {type: noise, scale: 0.04}}
  plain: {type: lambertian, color: [0.5, 0.5, 0.5]}
surfaces:
  - {type: mesh, vertices: [[0,0,0],[1,0,0],[0,0,1]], triangles: [[0,1,2]], material: floor}
  - {type: mesh, vertices: [[0,0,0],[1,0,0],[0,1,0]], triangles: [[0,1,2]], material: wall}
  - {type: mesh, vertices: [[0,1,0],[1,1,0],[0,1,1]], triangles: [[0,1,2]], material: lamp}
  - {type: mesh, vertices: [[0,0,0],[0,1,0],[0,0,1]], triangles: [[0,1,2]], material: marble}
  - {type: mesh, vertices: [[1,0,0],[1,1,0],[1,0,1]], triangles: [[0,1,2]], material: plain}
""")
    s = bdpt.SceneFile(str(tmp_path / "scene.yaml"))
    assert len(s.textures) == 4 and s.n_lights == 1
    mats = [s.material_list()[i] for i in s.mat_id]
    assert [len(m) for m in mats] == [4, 4, 4, 4, 3] and [m[3] for m in mats[:4]] == [1, 2, 3, 4]
    t = s.textures
    assert (t[0]["type"], t[0]["scale"], t[0]["a"], t[0]["b"]) == (bdpt.TEX_CHECKER, 60.0, (0.8, 0.8, 0.8), (0.15, 0.2, 0.6))
    assert (t[2]["type"], t[2]["a"], t[2]["b"]) == (bdpt.TEX_CHECKER, (15.0, 15.0, 15.0), (15.0, 6.0, 2.0))   # a light's colours stay linear HDR
    assert t[1]["type"] == bdpt.TEX_IMAGE and (t[1]["width"], t[1]["height"]) == (12, 9)
    assert np.array_equal(t[1]["data"], g["image_bytes"])
    assert t[3]["type"] == bdpt.TEX_NOISE and np.array_equal(t[3]["data"], g["noise_tables"])
