"""One single-threaded worker of the CPU reference arm (bench.py --impl reference / cpu_baseline).

Runs the reference's OWN code (oracle/_ref/libref_oracle.so: its unmodified headers behind
oracle/ref_harness.cpp) — or the oracle port when that library is absent — on one row of the workload
per step, and prints one JSON line. Independent single-thread processes over disjoint rows are the best
honest use of all host cores: the reference's own thread pool anti-scales.
usage: cpu_ref_worker.py scene.yaml W H spp depth row steps warmup [threads [rows]]
threads > 1: ONE process rendering `rows` rows per step through the reference's own thread pool shape (camera.h:60,
111-134: dynamic row counter, shared unlocked mt19937, shared stats atomics) -- what a user of the reference gets.
"""
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    scene, W, H, spp, depth, row, steps, warmup = sys.argv[1], *[int(x) for x in sys.argv[2:9]]
    threads = int(sys.argv[9]) if len(sys.argv) > 9 else 1
    rows = int(sys.argv[10]) if len(sys.argv) > 10 else 1
    from oracle.oracle import Oracle, available, Camera
    pkg = importlib.import_module("bidirectional-path-tracer_b200")
    sf = pkg.SceneFile(scene)                      # host-only loader; no GPU involved
    kind = "ref" if available("ref") else "port"
    t0 = time.time()
    o = Oracle(kind).load(sf.verts, sf.mat_id, sf.material_list())
    build_s = time.time() - t0
    cam = Camera.from_buffer_copy(bytes(sf.camera_with(W, H, spp, depth)))
    n = int(spp ** 0.5)
    times = []
    o.stats_reset()
    for s in range(warmup + steps):
        r = (row + 7 * s) % H                      # a different row every step
        t0 = time.time()
        r = min(r, H - rows)
        o.render(cam, row0=r, row1=r + rows, height=H, nthreads=threads)
        times.append(time.time() - t0)
    st = o.stats()
    print(json.dumps({"kind": "reference" if kind == "ref" else "port", "samples_per_step": W * n * n * rows, "threads": threads,
                      "step_s": times[warmup:], "build_s": build_s, "rays": st["rays"], "node_visits": st["node_visits"],
                      "tri_tests": st["tri_tests"], "all_steps": warmup + steps}))


if __name__ == "__main__":
    main()
