"""Scene image (include/bdpt.h "scene image", csrc/scene_image.cu): the prebuilt soup + both BVHs
as one block.  CPU tests: the build is deterministic, the header refuses truncated / damaged / foreign blocks, a
file round trip is byte-exact, and the YAML front-end's cache hits only while the sources are unchanged.
GPU tests: tracing and rendering from an image give exactly what bdpt_scene_upload of the same soup gives
(golden vectors of the reference included)."""
import os

import numpy as np
import pytest

from test_trace_gpu import NAMES, check_hits


def rel_err(a, b):
    return float(np.max(np.abs(a - b)) / max(1e-300, float(np.max(np.abs(b)))))


def soup(n=4000, seed=3):
    from tools import testscenes as ts
    return ts.random_soup(n, seed=seed, extent=10.0, size=0.6)


def test_image_build_is_deterministic_and_self_describing(bdpt, scenes):
    v, m, mats, cam = scenes["cornell_spheres"]()
    with bdpt.SceneImage.build(v, m, mats, source_key=0xABCDEF) as a, bdpt.SceneImage.build(v, m, mats, source_key=0xABCDEF) as b:
        ia = a.info(verify=True)
        assert ia["n_tris"] == len(v) and ia["source_key"] == 0xABCDEF and ia["n_mats"] == len(mats)
        order, n_ref, n_fast = bdpt.host_bvh_build(v, fast=True)
        assert ia["n_ref_nodes"] == n_ref and ia["n_fast_nodes"] == n_fast
        assert 0 < ia["device_bytes"] < a.nbytes and a.nbytes % 256 == 0
        assert np.array_equal(a.bytes(), b.bytes())          # threads inside the builders must not show in the bytes
        # the host-only tail of the image is the DFS leaf order of the reference topology
        tail = a.bytes()[512 + ia["device_bytes"]:][:4 * len(v)].view(np.uint32)
        assert np.array_equal(tail, order)


def test_image_reference_topology_matches_golden(bdpt, scenes, golden):
    for name in ("cornell_box", "unit_room"):
        g = golden(name)
        v, m, mats, cam = scenes[name]()
        with bdpt.SceneImage.build(v, m, mats) as img:
            i = img.info()
            assert i["n_ref_nodes"] == int(g["nodes"])
            tail = img.bytes()[512 + i["device_bytes"]:][:4 * len(v)].view(np.uint32)
            assert np.array_equal(tail, g["leaf_order"])


def image_sections(raw):
    """{index: bytes} from the image's section table (header: 128 bytes of fields, then {offset, bytes} pairs)."""
    n_sec = 15
    table = raw[128:128 + 16 * n_sec].view(np.uint64).reshape(n_sec, 2)
    return [raw[int(o):int(o) + int(b)] for o, b in table]


def test_wide_tree_is_the_binary_tree_collapsed(bdpt, scenes, monkeypatch):
    """The 4-wide nodes (fast_bvh.cu: collapse_wide, built when BDPT_WIDE asks for them) must hold exactly the binary
    fast tree's leaves, each once, under copied (never shrunk) boxes, with every wide node referenced once."""
    with bdpt.SceneImage.build(*soup(100, seed=8)) as img:
        assert len(image_sections(img.bytes())[6]) == 0                # the section is optional: empty by default
    monkeypatch.setenv("BDPT_WIDE", "1")
    for v, m, mats in (scenes["cornell_spheres"]()[:3], soup(5000, seed=4), soup(3, seed=1), soup(1, seed=2)):
        with bdpt.SceneImage.build(v, m, mats) as img:
            sec = image_sections(img.bytes())
        fast = sec[5].view(np.dtype([("box", "<f4", 12), ("left", "<i4"), ("right", "<i4"), ("pad", "<u4", 2)]))
        wide = sec[6].view(np.dtype([("box", "<f4", (6, 4)), ("child", "<i4", 4), ("pad", "<u4", 4)]))
        links = np.concatenate([fast["left"], fast["right"]])
        bin_leaves = np.sort(links[links < 0])
        ch = wide["child"].reshape(-1)
        empty = ch == np.iinfo(np.int32).min
        wide_leaves = np.sort(ch[(ch < 0) & ~empty])
        assert np.array_equal(bin_leaves, wide_leaves)
        inner = np.sort(ch[ch >= 0])
        assert np.array_equal(inner, np.arange(1, len(wide)))          # every node but the root has one parent
        assert len(wide) <= len(fast) and (len(fast) < 8 or len(wide) < 0.75 * len(fast))
        # every used child box of a wide node is one of the binary tree's child boxes
        bin_boxes = {tuple(b) for b in np.concatenate([fast["box"][:, [0, 1, 2, 3, 4, 5]], fast["box"][:, [6, 7, 8, 9, 10, 11]]]).tolist()}
        wb = wide["box"].transpose(0, 2, 1).reshape(-1, 6)[~empty]
        assert all(tuple(b) in bin_boxes for b in wb.tolist())


def test_image_rejects_damage(bdpt):
    v, m, mats = soup(500)
    with bdpt.SceneImage.build(v, m, mats) as img:
        good = img.bytes()
    assert bdpt.SceneImage.from_bytes(good).info(verify=True)["n_tris"] == 500
    with pytest.raises(bdpt.BdptError, match="truncated"):
        bdpt.SceneImage.from_bytes(good[:-256].copy()).info()
    with pytest.raises(bdpt.BdptError, match="shorter"):
        bdpt.SceneImage.from_bytes(good[:100].copy()).info()
    bad = good.copy(); bad[0] ^= 1
    with pytest.raises(bdpt.BdptError, match="magic"):
        bdpt.SceneImage.from_bytes(bad).info()
    bad = good.copy(); bad[8] = 99
    with pytest.raises(bdpt.BdptError, match="version"):
        bdpt.SceneImage.from_bytes(bad).info()
    bad = good.copy(); bad[len(bad) // 2] ^= 0x10            # one flipped payload bit
    assert bdpt.SceneImage.from_bytes(bad).info(verify=False)["n_tris"] == 500   # the header alone cannot see it
    with pytest.raises(bdpt.BdptError, match="hash"):
        bdpt.SceneImage.from_bytes(bad).info(verify=True)
    bad = good.copy(); bad[56:60] = np.frombuffer(np.uint32(501).tobytes(), np.uint8)   # n_tris no longer matches the table
    with pytest.raises(bdpt.BdptError, match="section table"):
        bdpt.SceneImage.from_bytes(bad).info()


def test_image_build_errors(bdpt):
    with pytest.raises(bdpt.BdptError):
        bdpt.SceneImage.build(np.zeros((0, 9)), np.zeros(0, np.uint32), [(0, (0.5, 0.5, 0.5), 0.0)])
    v, m, mats = soup(10)
    with pytest.raises(bdpt.BdptError):
        bdpt.SceneImage.build(v, m + 7, mats)                # material index out of range


def test_image_file_round_trip(bdpt, tmp_path):
    v, m, mats = soup(3000, seed=11)
    p = str(tmp_path / "scene.bdptimg")
    with bdpt.SceneImage.build(v, m, mats, source_key=5) as img:
        img.save(p)
        want = img.bytes()
    assert not os.path.exists(p + ".tmp")
    with bdpt.SceneImage.map(p) as mp:
        assert mp.info(verify=True)["source_key"] == 5
        assert np.array_equal(mp.bytes(), want)
    with pytest.raises(bdpt.BdptError):
        bdpt.SceneImage.map(str(tmp_path / "absent.bdptimg"))


YAML = """camera: {resolution: [8, 6], location: [0, 1, 5], look_at: [0, 1, 0], samples_per_pixel: 4, max_depth: 5, output: t.png}
materials:
  - {name: grey, type: lambertian, color: [0.5, 0.5, 0.5]}
surfaces:
  - type: object
    file: m.obj
    material: grey
  - type: mesh
    material: {type: light, color: [10, 10, 10]}
    vertices: [[-1, 3, -1], [1, 3, -1], [1, 3, 1], [-1, 3, 1]]
    triangles: [[0, 1, 2], [0, 2, 3]]
"""
OBJ = "v -2 0 -2\nv 2 0 -2\nv 2 0 2\nv -2 0 2\nf 1 2 3 4\n"


def test_yaml_cache_hits_only_while_sources_are_unchanged(bdpt, tmp_path):
    y, o, c = tmp_path / "s.yaml", tmp_path / "m.obj", str(tmp_path / "s.bdptimg")
    y.write_text(YAML)
    o.write_text(OBJ)
    plain = bdpt.SceneFile(str(y))
    with bdpt.SceneImage.build(plain.verts, plain.mat_id, plain.materials) as direct:
        want_payload = direct.bytes()[512:]
    key = bdpt.scene_source_key(str(y))
    with bdpt.CachedSceneFile(str(y), c) as f:
        assert not f.hit and f.n_tris == 4 and os.path.exists(c)
        assert f.image.info(verify=True)["source_key"] == key
        assert np.array_equal(f.image.bytes()[512:], want_payload)   # same soup -> same image as the direct path
        cam0 = bytes(f.camera)
    with bdpt.CachedSceneFile(str(y), c) as f:
        assert f.hit and f.n_tris == 4 and f.n_lights == 2 and f.output == "t.png"
        assert bytes(f.camera) == cam0 == bytes(plain.camera)
        assert np.array_equal(f.image.bytes()[512:], want_payload)
    o.write_text(OBJ.replace("v 2 0 2", "v 2 0 3"))             # geometry edited: the key changes, the cache is rebuilt
    assert bdpt.scene_source_key(str(y)) != key
    with bdpt.CachedSceneFile(str(y), c) as f:
        assert not f.hit
        assert not np.array_equal(f.image.bytes()[512:], want_payload)
    with bdpt.CachedSceneFile(str(y), c) as f:
        assert f.hit
    with open(c, "r+b") as fh:                                  # damaged cache file: rebuilt, not trusted
        fh.seek(4096); fh.write(b"\xff" * 8)
    with bdpt.CachedSceneFile(str(y), c) as f:
        assert not f.hit
    y.write_text(YAML.replace("max_depth: 5", "max_depth: 7"))  # camera-only edit also changes the key
    with bdpt.CachedSceneFile(str(y), c) as f:
        assert not f.hit and f.camera.max_depth == 7
    with pytest.raises(bdpt.BdptError, match="Failed to load scene"):
        bdpt.CachedSceneFile(str(tmp_path / "absent.yaml"), c)


# ---------------------------------------------------------------------------------------------- GPU parity
@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_image_upload_trace_matches_reference_golden(name, bdpt, gpu_ctx, scenes, golden):
    g = golden(name)
    v, m, mats, cam = scenes[name]()
    with bdpt.SceneImage.build(v, m, mats) as img:
        gpu_ctx.upload_image(img, verify=True)
    assert gpu_ctx.bvh_info(bdpt.BVH_REFERENCE)[0] == int(g["nodes"])
    assert np.array_equal(gpu_ctx.leaf_order(), g["leaf_order"])
    for bvh in (bdpt.BVH_REFERENCE, bdpt.BVH_FAST):
        h = gpu_ctx.trace(g["rays"], bvh=bvh)
        check_hits(h, g["prim"], g["t"], g["u"], g["v"], exact=True)
        assert np.array_equal(gpu_ctx.visible(g["seg_a"], g["seg_b"], bvh=bvh), g["visible"])


@pytest.mark.gpu
def test_image_upload_renders_the_same_film(bdpt, gpu_ctx, scenes, tmp_path):
    v, m, mats, cam = scenes["cornell_spheres"]()
    c = bdpt.make_camera(48, 40, 16, 8, **cam)
    gpu_ctx.upload(v, m, mats)
    gpu_ctx.render(c, seed=3, bvh=bdpt.BVH_FAST, height=40)
    want = gpu_ctx.film()
    p = str(tmp_path / "c.bdptimg")
    with bdpt.SceneImage.build(v, m, mats) as img:
        img.save(p)
    with bdpt.SceneImage.map(p) as img:
        gpu_ctx.upload_image(img, verify=True)
    gpu_ctx.render(c, seed=3, bvh=bdpt.BVH_FAST, height=40)
    assert rel_err(gpu_ctx.film(), want) < 1e-12                # fp64 atomics: summation order only
    gpu_ctx.upload(v, m, mats)                                  # and back: an image upload leaves no state behind
    gpu_ctx.render(c, seed=3, bvh=bdpt.BVH_REFERENCE, height=40)
    ref_film = gpu_ctx.film()
    with bdpt.SceneImage.map(p) as img:
        gpu_ctx.upload_image(img)
    gpu_ctx.render(c, seed=3, bvh=bdpt.BVH_REFERENCE, height=40)
    assert rel_err(gpu_ctx.film(), ref_film) < 1e-12


def corrupt_fast_link(good):
    """One FastNode child link sent far outside the tree; the payload hash is left stale (a caller that does not ask for
    the hash check must still be stopped by the range check)."""
    import struct
    bad = good.copy()
    off = struct.unpack_from("<Q", bad, 128 + 16 * 5)[0]        # section table (header offset 128), SEC_FAST_NODES = 5
    bad[off + 48:off + 52] = np.frombuffer(np.int32(0x7FFFFF00).tobytes(), np.uint8)   # node 0, left link (after 12 floats)
    return bad


def test_image_range_check_catches_wild_links(bdpt):
    """links / prim / material / light ids of the payload are range-checked (bdpt_scene_image_info verify & 2,
    and every bdpt_scene_upload_image without BDPT_IMAGE_TRUSTED)."""
    import struct
    v, m, mats = soup(500)
    with bdpt.SceneImage.build(v, m, mats) as img:
        good = img.bytes()
        assert img.info(verify=True, validate=True)["n_tris"] == 500
    bad = corrupt_fast_link(good)
    assert bdpt.SceneImage.from_bytes(bad).info()["n_tris"] == 500          # header and section table are intact
    with pytest.raises(bdpt.BdptError, match="fast-tree link out of range"):
        bdpt.SceneImage.from_bytes(bad).info(validate=True)
    bad = good.copy()
    off = struct.unpack_from("<Q", bad, 128 + 16 * 0)[0]                    # SEC_TRIS: TriGeom.mat at byte 72
    bad[off + 72:off + 76] = np.frombuffer(np.uint32(77).tobytes(), np.uint8)
    with pytest.raises(bdpt.BdptError, match="material id out of range"):
        bdpt.SceneImage.from_bytes(bad).info(validate=True)
    bad = good.copy()
    off = struct.unpack_from("<Q", bad, 128 + 16 * 3)[0]                    # SEC_LIGHTS
    bad[off:off + 4] = np.frombuffer(np.uint32(500).tobytes(), np.uint8)
    with pytest.raises(bdpt.BdptError, match="light id out of range"):
        bdpt.SceneImage.from_bytes(bad).info(validate=True)


def test_concurrent_savers_do_not_tear_the_cache_file(bdpt, tmp_path):
    """several ranks missing the same cache at once each write a temporary file of their own."""
    import threading
    v, m, mats = soup(2000)
    path = str(tmp_path / "shared.bdptimg")
    with bdpt.SceneImage.build(v, m, mats) as img:
        good = img.bytes()
        errs = []
        def save():
            try:
                for _ in range(5):
                    img.save(path)
                    with bdpt.SceneImage.map(path) as mp:
                        assert mp.info(verify=True)["n_tris"] == 2000
            except Exception as e:   # noqa: BLE001
                errs.append(e)
        ts = [threading.Thread(target=save) for _ in range(6)]
        [t.start() for t in ts]; [t.join() for t in ts]
    assert not errs
    assert sorted(os.listdir(tmp_path)) == ["shared.bdptimg"]               # no temporary file left behind
    assert np.array_equal(np.fromfile(path, np.uint8), good)


@pytest.mark.gpu
def test_image_upload_refuses_wild_links_and_keeps_the_old_scene(bdpt, gpu_ctx, scenes, golden):
    g = golden("cornell_box")
    v, m, mats, cam = scenes["cornell_box"]()
    gpu_ctx.upload(v, m, mats)
    with bdpt.SceneImage.build(v, m, mats) as img:
        bad = corrupt_fast_link(img.bytes())
    with pytest.raises(bdpt.BdptError, match="out of range"):
        gpu_ctx.upload_image(bdpt.SceneImage.from_bytes(bad), verify=False)   # no hash check asked for: the range check stops it
    h = gpu_ctx.trace(g["rays"], bvh=bdpt.BVH_FAST)
    check_hits(h, g["prim"], g["t"], g["u"], g["v"], exact=True)


@pytest.mark.gpu
def test_image_upload_refuses_damage_and_keeps_the_old_scene(bdpt, gpu_ctx, scenes, golden):
    g = golden("cornell_box")
    v, m, mats, cam = scenes["cornell_box"]()
    gpu_ctx.upload(v, m, mats)
    with bdpt.SceneImage.build(v, m, mats) as img:
        bad = img.bytes()
    bad[len(bad) // 2] ^= 1
    with pytest.raises(bdpt.BdptError, match="hash"):
        gpu_ctx.upload_image(bdpt.SceneImage.from_bytes(bad), verify=True)
    h = gpu_ctx.trace(g["rays"], bvh=bdpt.BVH_FAST)             # the refused upload did not touch the resident scene
    check_hits(h, g["prim"], g["t"], g["u"], g["v"], exact=True)


@pytest.mark.gpu
def test_cli_cache_renders_the_same_png(bdpt, tmp_path):
    """`bdpt scene.yaml --cache f`: the first run builds and writes the image, the second maps it (no OBJ parse, no
    tree build); both PNGs must equal the uncached run's within the film's atomic summation order."""
    import subprocess
    from test_loader import decode_png
    from tools import make_standin_scenes as ms
    ms.generate(detail=2, quiet=True)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "bidirectional-path-tracer_b200", "bdpt")
    scene = os.path.join(root, "scenes", "coffee_machine.yaml")
    base = [exe, scene, "--res", "64x36", "--spp", "4", "--depth", "6", "--seed", "3"]
    cache = str(tmp_path / "coffee.bdptimg")
    pngs = []
    for k, extra in enumerate(([], ["--cache", cache], ["--cache", cache])):
        r = subprocess.run(base + ["--out", "c%d.png" % k] + extra, cwd=tmp_path, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-1000:]
        if k == 1:
            assert "built and written" in r.stderr and "Triangles:" in r.stdout
        if k == 2:
            assert "cache hit" in r.stderr and "Triangles:" not in r.stdout    # the OBJ files were not parsed
        pngs.append(decode_png(tmp_path / "output" / ("c%d.png" % k)))
    for p in pngs[1:]:
        d = np.abs(p.astype(int) - pngs[0].astype(int))
        assert d.max() <= 1 and (d > 0).mean() < 0.01


@pytest.mark.gpu
def test_cli_two_gpus_share_one_scene_image(bdpt, tmp_path):
    """`bdpt --gpus 2`: camera::render builds the scene image once and uploads it to both contexts (host/camera.cpp);
    strata are sharded, the films summed by ONE NCCL reduce inside the library (bdpt_film_reduce) -- the PNG must equal
    the one-GPU run's, and the run must say that the reduce happened."""
    import subprocess
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from test_loader import decode_png
    from tools import make_standin_scenes as ms
    ms.generate(detail=2, quiet=True)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "bidirectional-path-tracer_b200", "bdpt")
    scene = os.path.join(root, "scenes", "coffee_machine.yaml")
    pngs = []
    for g in (1, 2):
        r = subprocess.run([exe, scene, "--res", "96x54", "--spp", "16", "--depth", "8", "--seed", "4", "--gpus", str(g), "--out", "g%d.png" % g],
                           cwd=tmp_path, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stderr[-1000:]
        assert "on %d GPU(s)" % g in r.stderr
        assert ("film reduce (NCCL)" in r.stderr) == (g == 2)
        assert "BVH Stats:" in r.stderr and "Rays traced:" in r.stderr      # stats.h:34-50 at the end of camera::render
        pngs.append(decode_png(tmp_path / "output" / ("g%d.png" % g)))
    d = np.abs(pngs[0].astype(int) - pngs[1].astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 0.01


@pytest.mark.gpu
def test_film_reduce_sums_the_shards_on_two_gpus(bdpt, scenes):
    """bdpt_film_reduce (one process, two devices): each context renders its strata, one NCCL sum-reduce onto context 0;
    the result equals the single-GPU film up to fp64 summation order."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    v, m, mats, cam = scenes["cornell_spheres"]()
    c = bdpt.make_camera(48, 32, 16, 6, **cam)
    with bdpt.Context(0) as a, bdpt.Context(1) as b:
        for ctx in (a, b):
            ctx.upload(v, m, mats)
        a.render(c, seed=9, bvh=bdpt.BVH_FAST, height=32)
        whole = a.film().copy()
        a.render(c, seed=9, bvh=bdpt.BVH_FAST, height=32, strata=bdpt.shard_strata(0, 2, 16))
        b.render(c, seed=9, bvh=bdpt.BVH_FAST, height=32, strata=bdpt.shard_strata(1, 2, 16))
        ms = bdpt.film_reduce([a, b], root=0)
        assert ms >= 0
        assert np.abs(a.film() - whole).max() / np.abs(whole).max() < 1e-12
