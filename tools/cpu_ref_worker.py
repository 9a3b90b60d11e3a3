"""One single-threaded worker of the CPU reference arm (bench.py --impl reference / cpu_baseline).

Runs the reference's OWN code (oracle/_ref/libref_oracle.so: its unmodified headers behind
oracle/ref_harness.cpp) — or the oracle port when that library is absent — and prints one JSON line.  The scene comes
from an .npz written by tools/dump_soup.py, so this process maps neither libbdpt.so nor CUDA.  Independent
single-thread processes are the best honest use of all host cores: the reference's own thread pool anti-scales
(SURVEY F5).  Each step renders `rows` rows spread evenly over the image height, offset by this worker's index, so
all workers see the same mix of cheap (background) and dear (object) rows.
usage: cpu_ref_worker.py soup.npz W H spp depth worker n_workers steps warmup [threads [rows]]
threads > 1: ONE process rendering `rows` consecutive rows per step through the reference's own thread pool shape
(camera.h:60, 111-134: dynamic row counter, shared unlocked mt19937, shared stats atomics) -- what a user of the
reference gets.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    soup = sys.argv[1]
    W, H, spp, depth, worker, n_workers, steps, warmup = [int(x) for x in sys.argv[2:10]]
    threads = int(sys.argv[10]) if len(sys.argv) > 10 else 1
    rows = int(sys.argv[11]) if len(sys.argv) > 11 else 1
    from oracle.oracle import Oracle, available, make_camera
    z = np.load(soup)
    mats = [(int(t), tuple(rgb), float(p)) for t, rgb, p in zip(z["mat_type"], z["mat_rgb"], z["mat_param"])]
    kind = "ref" if available("ref") else "port"
    t0 = time.time()
    o = Oracle(kind).load(z["verts"], z["mat_id"], mats)
    build_s = time.time() - t0
    c = z["cam"]
    cam = make_camera(W, H, spp, depth, lookfrom=tuple(c[1:4]), lookat=tuple(c[4:7]), vfov=float(c[0]), vup=tuple(c[7:10]),
                      background=tuple(c[10:13]), focus_dist=float(c[13]))
    n = int(spp ** 0.5)
    times = []
    o.stats_reset()
    for s in range(warmup + steps):
        t0 = time.time()
        if threads > 1:
            r = min((H // 2 + 7 * s) % H, H - rows)
            o.render(cam, row0=r, row1=r + rows, height=H, nthreads=threads)
        else:
            for k in range(rows):   # rows spread over the height, interleaved between the workers and shifted every step
                r = (int((k * n_workers + worker + 0.5) * H / (rows * n_workers)) + 7 * s) % H
                o.render(cam, row0=r, row1=r + 1, height=H, nthreads=1)
        times.append(time.time() - t0)
    st = o.stats()
    print(json.dumps({"kind": "reference" if kind == "ref" else "port", "samples_per_step": W * n * n * rows, "threads": threads,
                      "step_s": times[warmup:], "build_s": build_s, "rays": st["rays"], "node_visits": st["node_visits"],
                      "tri_tests": st["tri_tests"], "all_steps": warmup + steps}))


if __name__ == "__main__":
    main()
