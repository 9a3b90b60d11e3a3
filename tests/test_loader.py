"""Host front-end (CPU only): the YAML/OBJ loader against a pyyaml cross-check and the PNG
writer against a decoder (T10). The reference has no tests for either; its loader behaviour is restated
from src/scene/scene_loader.h and pinned here."""
import os
import shutil
import zlib

import numpy as np
import pytest
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def standins():
    from tools import make_standin_scenes as ms
    ms.generate(detail=1 if os.environ.get("BDPT_FAST_TESTS") else 2, quiet=True)
    return os.path.join(ROOT, "scenes")


def obj_faces(path):
    n = 0
    with open(path) as fh:
        for line in fh:
            if line.startswith("f "):
                n += len(line.split()) - 3
    return n


@pytest.mark.parametrize("name", ["coffee_machine", "glass-of-water"])
def test_named_scene_matches_pyyaml_view(name, bdpt, standins):
    path = os.path.join(standins, name + ".yaml")
    doc = yaml.safe_load(open(path, encoding="utf-8"))
    s = bdpt.SceneFile(path)
    cam = doc["camera"]
    W, H = cam["resolution"]
    assert s.camera.image_width == W and s.camera.aspect_ratio == W / H
    assert s.camera.vfov == float(cam.get("vfov", cam.get("fov", 90)))
    assert list(s.camera.lookfrom) == cam["location"] and list(s.camera.lookat) == cam["look_at"]
    assert s.camera.samples_per_pixel == cam.get("samples_per_pixel", 50)      # camera.h:28 default
    assert s.camera.max_depth == cam.get("max_depth", 10)                        # camera.h:29 default
    assert s.camera.defocus_angle == 0.0 and s.camera.focus_dist == 10.0
    assert s.output == cam.get("output", "image.png")
    # triangles: OBJ faces (fan-triangulated) + inline meshes, in file order
    expect, lights = 0, 0
    for surf in doc["surfaces"]:
        if surf["type"] == "object":
            expect += obj_faces(os.path.join(standins, surf["file"]))
        elif surf["type"] == "mesh":
            expect += len(surf["triangles"])
            if doc["materials"][surf["material"]]["type"] == "light":
                lights += len(surf["triangles"])
    assert len(s.verts) == expect and s.n_lights == lights
    # material table: 0-255 colours scaled, emission kept linear (scene_loader.h:81-92,123-125)
    mats = s.material_list()
    for t, rgb, p in mats:
        if t == bdpt.MAT_LIGHT:
            assert max(rgb) > 1.0
        elif t in (bdpt.MAT_LAMBERTIAN, bdpt.MAT_METAL):
            assert 0.0 <= min(rgb) and max(rgb) <= 1.0
    light_tris = s.verts[np.array([mats[i][0] for i in s.mat_id]) == bdpt.MAT_LIGHT]
    assert len(light_tris) == lights
    # the inline light quad vertices arrive bit-exactly
    last = [x for x in doc["surfaces"] if x["type"] == "mesh"][-1]
    v, t = np.array(last["vertices"], dtype=np.float64), last["triangles"]
    assert np.array_equal(s.verts[-1], np.concatenate([v[t[-1][0]], v[t[-1][1]], v[t[-1][2]]]))


@pytest.mark.skipif(not os.path.exists("/root/reference/scenes"), reason="reference tree absent (GPU box)")
@pytest.mark.parametrize("name", ["coffee_machine", "glass-of-water"])
def test_reference_yaml_files_load_identically(name, bdpt, standins, tmp_path):
    """The reference's own YAML text (read here only, never on the GPU box) over the stand-in OBJs gives
    the same scene as the restated scenes/*.yaml."""
    shutil.copy("/root/reference/scenes/%s.yaml" % name, tmp_path / (name + ".yaml"))
    os.symlink(os.path.join(standins, "data"), tmp_path / "data")
    a = bdpt.SceneFile(str(tmp_path / (name + ".yaml")))
    b = bdpt.SceneFile(os.path.join(standins, name + ".yaml"))
    assert np.array_equal(a.verts, b.verts) and np.array_equal(a.mat_id, b.mat_id)
    assert a.material_list() == b.material_list()
    assert bytes(a.camera) == bytes(b.camera) and a.output == b.output


SYNTH = """
# comment with non-ascii: ≈ é
camera:
  resolution: [64, 32]   # trailing comment
  vfov: 200              # clamped to 179
  location: [0, 1, 5]
  look_at: [0, 0, 0]
  up: [0, 1, 0]
  background: [0.1, 0.2, 0.3]
  samples_per_pixel: 17
  max_depth: 7
  focus_distance: 3.5
  aperture_radius: 4
  output: "synth.png"
materials:
  red: {type: lambertian, albedo: [255, 0, 0]}
  broken: 12
  mirror:
    type: metal
    base_color: [0.5, 0.5, 0.5]
    roughness: 7
  legacy_glass:
    transmission: 1.0
    ior: -2
  legacy_light:
    emission: [100, 50, 0]
scene:
  - type: TriMesh
    data:
      vertices: [0,0,0, 1,0,0, 0,1,0,
                 0,0,1, 1,0,1, 0,1,1]
    material:
      type: diffuse_light
      emission: [300, 300, 300]
  - type: Sphere
    data: {center: [0, 0, -2], radius: 0.5}
    material: {type: glass}
  - type: mesh
    vertices: [[0,0,0],[1,0,0],[1,1,0],[0,1,0]]
    triangles: [[0,1,2],[0,2,3],[0,1]]
    material: mirror
  - type: mesh
    vertices: [[0,0,5],[1,0,5],[1,1,5]]
    triangles: [[0,1,2]]
    material: nosuch
  - type: mesh
    vertices: [[0,0,6],[1,0,6],[1,1,6]]
    triangles: [[0,1,2]]
    material: legacy_light
  - type: Blob
"""


def test_schema_corners(bdpt, tmp_path):
    p = tmp_path / "s.yaml"
    p.write_text(SYNTH, encoding="utf-8")
    s = bdpt.SceneFile(str(p))
    c = s.camera
    assert (c.image_width, c.aspect_ratio, c.vfov, c.samples_per_pixel, c.max_depth) == (64, 2.0, 179.0, 17, 7)
    assert c.focus_dist == 3.5 and c.defocus_angle == 0.0 and list(c.background) == [0.1, 0.2, 0.3]
    assert s.output == "synth.png"
    assert len(s.verts) == 2 + 960 + 2 + 1 + 1            # Sphere -> 16x32 UV sphere = 960 triangles
    mats = [s.material_list()[i] for i in s.mat_id]
    assert mats[0] == (3, (300.0, 300.0, 300.0), 0.0)     # emission is never divided by 255
    assert mats[2] == (2, (1.0, 1.0, 1.0), 1.5)           # glass default ior
    assert mats[962] == (1, (0.5, 0.5, 0.5), 1.0)         # roughness clamped to [0,1]
    assert mats[964] == (0, (0.8, 0.8, 0.8), 0.0)         # unknown material name -> grey lambertian
    assert mats[965] == (3, (100 / 255.0, 50 / 255.0, 0.0), 0.0)   # legacy emission in (1,255] IS scaled (scene_loader.h:143)
    assert s.n_lights == 3


@pytest.mark.parametrize("text,msg", [
    ("camera: {fov: 30}\nsurfaces: []\n", "resolution"),
    ("camera: {resolution: [0, 5]}\nsurfaces: []\n", "positive"),
    ("camera: {resolution: [4, 4]}\n", "missing"),
    ("camera: {resolution: [4, 4]}\nsurfaces:\n  - type: object\n    file: nope.obj\n", "OBJ"),
    ("camera: {resolution: [4, 4]}\nsurfaces:\n  - type: mesh\n    vertices: [[0,0,0]]\n    triangles: [[0,1,2]]\n", ""),
    ("- 1\n- 2\n", "mapping"),
])
def test_loader_errors(bdpt, tmp_path, text, msg):
    p = tmp_path / "bad.yaml"
    p.write_text(text)
    with pytest.raises(bdpt.BdptError, match="Failed to load scene.*" + msg):
        bdpt.SceneFile(str(p))


def test_obj_reader_indices_and_fans(bdpt, tmp_path):
    (tmp_path / "m.obj").write_text("# c\nv 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvn 0 0 1\nvt 0 0\n"
                                    "f 1/1/1 2/1/1 3/1/1 4/1/1\nf -4//1 -3//1 -2//1\nf 1 2\n")
    (tmp_path / "s.yaml").write_text("camera: {resolution: [4, 4]}\nsurfaces:\n  - type: object\n    file: m.obj\n    material: {type: lambertian, color: [0.5,0.5,0.5]}\n")
    s = bdpt.SceneFile(str(tmp_path / "s.yaml"))
    assert len(s.verts) == 3                                # quad -> 2 (fan), negative-index tri -> 1, 2-vertex face dropped
    assert s.verts[1].tolist() == [0, 0, 0, 1, 1, 0, 0, 1, 0]
    assert s.verts[2].tolist() == [0, 0, 0, 1, 0, 0, 1, 1, 0]


def decode_png(path):
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, W, H = 8, b"", 0, 0
    while pos < len(data):
        n = int.from_bytes(data[pos:pos + 4], "big")
        typ, body = data[pos + 4:pos + 8], data[pos + 8:pos + 8 + n]
        assert zlib.crc32(typ + body) == int.from_bytes(data[pos + 8 + n:pos + 12 + n], "big")
        if typ == b"IHDR":
            W, H = int.from_bytes(body[:4], "big"), int.from_bytes(body[4:8], "big")
            assert body[8:] == bytes([8, 2, 0, 0, 0])
        elif typ == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(H, W * 3 + 1)
    assert not raw[:, 0].any()
    return raw[:, 1:].reshape(H, W, 3)


def test_png_round_trip(bdpt, tmp_path):
    rgb = np.random.default_rng(0).integers(0, 256, (13, 7, 3), dtype=np.uint8)
    bdpt.write_png(str(tmp_path / "x.png"), rgb)
    assert np.array_equal(decode_png(tmp_path / "x.png"), rgb)
    try:
        from PIL import Image
        assert np.array_equal(np.asarray(Image.open(tmp_path / "x.png")), rgb)
    except ImportError:
        pass
