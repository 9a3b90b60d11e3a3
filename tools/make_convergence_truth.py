#!/usr/bin/env python
"""CPU ground truth for the relMSE convergence tests (tests/test_convergence_gpu.py), rendered ONCE by the
reference's own code (oracle/_ref: its unmodified headers behind oracle/ref_harness.cpp) and committed as small
fixtures under tests/golden/.  TEST INFRASTRUCTURE: needs /root/reference (this container), not the GPU box.

  python tools/make_convergence_truth.py [scene ...]        # default: glass-of-water coffee_machine

Each named scene (stand-in geometry of tools/make_standin_scenes.py, detail 2) is rendered at a reduced resolution
with TRUTH_SPP samples per pixel (>= 64x the largest spp the test compares), one single-thread process per row,
every process at its own place of the reference's mt19937 stream (ref_render's seed).  The fixture holds the linear
AVERAGED film as float32 plus everything needed to re-create the camera.
"""
import importlib
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

# name -> (yaml, W, H, depth; 0 = the YAML's own max_depth); config 1 is glass-of-water (BASELINE.json: square, the YAML's
# depth 80), configs 3/4 coffee_machine at depth 16
SCENES = {"glass-of-water": ("scenes/glass-of-water.yaml", 64, 64, 0),
          "coffee_machine": ("scenes/coffee_machine.yaml", 96, 54, 16)}
TRUTH_SPP = 4096


def worker(yaml, W, H, depth, spp, row0, row1, seed, out):
    from oracle.oracle import Oracle, Camera
    pkg = importlib.import_module("bidirectional-path-tracer_b200")   # host-only YAML loader (the reference's needs yaml-cpp)
    sf = pkg.SceneFile(os.path.join(ROOT, yaml))
    o = Oracle("ref").load(sf.verts, sf.mat_id, sf.material_list())
    cam = Camera.from_buffer_copy(bytes(sf.camera_with(W, H, spp, depth if depth > 0 else None)))
    film = o.render(cam, row0=row0, row1=row1, seed=seed, height=H)
    np.save(out, film[row0:row1])


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--worker":
        a = sys.argv[2:]
        return worker(a[0], int(a[1]), int(a[2]), int(a[3]), int(a[4]), int(a[5]), int(a[6]), int(a[7]), a[8])
    from tools import make_standin_scenes as ms
    ms.generate(detail=2, quiet=True)
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s"])
    names = sys.argv[1:] or list(SCENES)
    procs = os.cpu_count() or 8
    for name in names:
        yaml, W, H, depth = SCENES[name]
        t0 = time.time()
        bands = [(r, min(H, r + 1)) for r in range(H)]           # one row per job: balances the load
        jobs, running, films = list(enumerate(bands)), [], {}
        tmp = os.path.join("/tmp", "truth_%s_%d" % (name, os.getpid()))
        os.makedirs(tmp, exist_ok=True)
        while jobs or running:
            while jobs and len(running) < procs:
                k, (r0, r1) = jobs.pop(0)
                out = os.path.join(tmp, "band%04d.npy" % k)
                cmd = [sys.executable, os.path.abspath(__file__), "--worker", yaml, str(W), str(H), str(depth), str(TRUTH_SPP), str(r0), str(r1),
                       str(1 + k), out]
                running.append((k, r0, r1, out, subprocess.Popen(cmd, stdout=subprocess.DEVNULL)))
            time.sleep(0.2)
            for job in list(running):
                if job[4].poll() is not None:
                    if job[4].returncode:
                        raise SystemExit("worker for rows %d-%d failed" % (job[1], job[2]))
                    films[job[0]] = np.load(job[3])
                    running.remove(job)
        film = np.concatenate([films[k] for k in sorted(films)], axis=0) / float(TRUTH_SPP)
        path = os.path.join(ROOT, "tests", "golden", "truth_%s.npz" % name.replace("-", "_"))
        meta = {"scene": yaml, "width": W, "height": H, "depth": depth, "spp": TRUTH_SPP, "renderer": "oracle/_ref (reference headers), mt19937, "
                "one process per row", "standin_detail": 2, "seconds": time.time() - t0, "host_cores": procs}
        np.savez_compressed(path, film=film.astype(np.float32), meta=json.dumps(meta))
        print(name, film.shape, "mean", float(film.mean()), "%.0f s" % (time.time() - t0), "->", path, flush=True)


if __name__ == "__main__":
    main()
