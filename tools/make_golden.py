"""Generate tests/golden/*.npz from the REFERENCE's own headers (oracle/_ref/libref_oracle.so).

Run in the build container (where /root/reference exists):  python tools/make_golden.py
Each fixture stores the seeded inputs' parameters and the reference outputs, so that the oracle port
and the CUDA path can be checked on the GPU box, where /root/reference does not exist.
The reference's mt19937 stream is process-global, so every render fixture is produced in a fresh
subprocess (stream restarts at the default seed, exactly like a fresh run of the reference binary).
"""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.oracle import Oracle, make_camera  # noqa: E402
from tools import testscenes as ts  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

SCENES = {
    "cornell_spheres": lambda: ts.cornell(True, (8, 12)),
    "cornell_box": lambda: ts.cornell(False),
    "unit_room": lambda: ts.unit_room(seed=1, n_blobs=6),
}
RENDERS = {  # name -> (scene, W, H, spp, depth)
    "cornell_spheres": (24, 24, 4, 8),
    "cornell_box": (16, 16, 4, 5),
    "unit_room": (24, 16, 9, 12),
}


def segments(verts, n, seed):
    """unoccluded-looking floor<->light style segments: random surface point pairs (fp32-representable)."""
    rays = ts.ray_batch(verts, 3 * n, seed=seed)
    seg = rays[-n:]
    a = seg[:, 0:3]
    b = a + seg[:, 3:6] * (seg[:, 7:8] + 0.001)
    return np.ascontiguousarray(a), np.ascontiguousarray(b)


def render_child(name):
    v, m, mats, cam = SCENES[name]()
    W, H, spp, depth = RENDERS[name]
    o = Oracle("ref").load(v, m, mats)
    film = o.render(make_camera(W, H, spp, depth, **cam), height=H)
    np.save(os.path.join(GOLD, "_tmp_film.npy"), film)


TEXTURED_RENDER = (20, 20, 4, 6)


def textured_child():
    """cornell_textured WITHOUT the noise box, so that the render starts at the head of the reference's mt19937 stream
    (a noise_texture draws its tables from it when it is built, perlin.h:6-14)."""
    from oracle import oracle as O
    v, m, mats, cam, texs = ts.cornell_textured(noise=False)
    tarr, blob = O.make_textures(texs)
    o = O.Oracle("ref").load(v, m, mats, textures=(tarr, len(texs), blob))
    W, H, spp, depth = TEXTURED_RENDER
    np.save(os.path.join(GOLD, "_tmp_film.npy"), o.render(make_camera(W, H, spp, depth, **cam), height=H))


def textured_fixture():
    """What the reference's own texture objects hold and return: the image bytes after rtw_image's float round trip,
    a noise texture's perlin tables, texture::value at 4,000 points for each texture, and a whole render."""
    from oracle import oracle as O
    v, m, mats, cam, texs = ts.cornell_textured(noise=True)
    tarr, blob = O.make_textures(texs)
    o = O.Oracle("ref").load(v, m, mats, textures=(tarr, len(texs), blob))
    image_bytes = o.texture_data(1, 12 * 9 * 3)
    noise_tables = o.texture_data(3, O.NOISE_BYTES)
    rng = np.random.default_rng(5)
    uvp = np.concatenate([rng.uniform(-0.2, 1.2, (4000, 2)), rng.uniform(-50, 600, (4000, 3))], axis=1)
    values = np.stack([o.texture_value(k, uvp) for k in range(len(texs))])
    subprocess.check_call([sys.executable, os.path.abspath(__file__), "--textured-child"])
    film = np.load(os.path.join(GOLD, "_tmp_film.npy"))
    os.remove(os.path.join(GOLD, "_tmp_film.npy"))
    np.savez_compressed(os.path.join(GOLD, "cornell_textured.npz"), image_bytes=image_bytes, noise_tables=noise_tables, uvp=uvp,
                        values=values, film=film, render=np.array(TEXTURED_RENDER))
    print("cornell_textured", len(v), "tris, film mean", film.mean(), "image bytes changed by the loader:", (image_bytes != texs[1]["data"].reshape(-1)).mean())


def main():
    if len(sys.argv) == 3 and sys.argv[1] == "--render-child":
        render_child(sys.argv[2])
        return
    if len(sys.argv) == 2 and sys.argv[1] == "--textured-child":
        textured_child()
        return
    if len(sys.argv) == 2 and sys.argv[1] == "--textured":
        textured_fixture()
        return
    os.makedirs(GOLD, exist_ok=True)
    textured_fixture()
    for name, mk in SCENES.items():
        v, m, mats, cam = mk()
        o = Oracle("ref").load(v, m, mats)
        rays = ts.ray_batch(v, 3000, seed=0x5EED0001, cam=cam)
        prim, t, u, w = o.trace(rays)
        a, b = segments(v, 1500, seed=77)
        vis = o.visible(a, b)
        subprocess.check_call([sys.executable, os.path.abspath(__file__), "--render-child", name])
        film = np.load(os.path.join(GOLD, "_tmp_film.npy"))
        os.remove(os.path.join(GOLD, "_tmp_film.npy"))
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), n_tris=len(v), nodes=o.node_count(),
                            leaf_order=o.leaf_order(), rays=rays, prim=prim, t=t, u=u, v=w, seg_a=a, seg_b=b,
                            visible=vis, film=film, render=np.array(RENDERS[name]))
        print(name, len(v), "tris", o.node_count(), "nodes, hit rate", (prim != 0xFFFFFFFF).mean(),
              "visible", vis.mean(), "film mean", film.mean())


if __name__ == "__main__":
    main()
