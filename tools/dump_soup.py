"""Flatten a YAML scene into an .npz (triangle soup, material table, camera fields) with the host-only loader of the
product library.  Runs in its OWN short-lived process: the CPU reference arm (bench.py --impl reference,
tools/cpu_ref_worker.py) then loads the .npz and never maps libbdpt.so (the reference's own scene_loader.h needs
yaml-cpp, which this image lacks: SURVEY F2).
usage: dump_soup.py scene.yaml out.npz"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    scene, out = sys.argv[1], sys.argv[2]
    pkg = importlib.import_module("bidirectional-path-tracer_b200")
    sf = pkg.SceneFile(scene)
    mats = sf.material_list()
    c = sf.camera
    np.savez(out, verts=sf.verts, mat_id=sf.mat_id, mat_type=np.array([m[0] for m in mats], np.uint32),
             mat_rgb=np.array([m[1] for m in mats], np.float64), mat_param=np.array([m[2] for m in mats], np.float64),
             cam=np.array([c.vfov, *c.lookfrom, *c.lookat, *c.vup, *c.background, c.focus_dist], np.float64))


if __name__ == "__main__":
    main()
