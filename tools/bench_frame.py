#!/usr/bin/env python
"""One WHOLE frame of a named scene, stratum-sharded over the ranks of a torchrun launch with one NCCL film
reduce — BASELINE.json config 4 by default: scenes/coffee_machine.yaml at 3840x2160, 1024 spp
(32x32 strata), max depth 16.  Total work is fixed, so running it at 1 and at N GPUs gives the STRONG-scaling
efficiency T1 / (N * TN) the north star asks for (bench.py's --gpus N is weak scaling).  Prints one JSON line on
rank 0 (not the bench.py contract; recorded in BASELINE.md).

  python tools/bench_frame.py [--res 3840x2160] [--spp 1024] [--depth 16] [--png out.png]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 tools/bench_frame.py

Timed region (max over ranks, barrier + synchronize on both sides): scene image upload (one H2D copy of the
prebuilt image), this rank's strata of every pixel, the film reduce onto rank 0, rank 0's film read-back.
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default=os.path.join(ROOT, "scenes", "coffee_machine.yaml"))
    ap.add_argument("--res", default="3840x2160")
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--depth", type=int, default=16)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--png", default="")
    args = ap.parse_args()
    import torch
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    if not torch.cuda.is_available():
        raise SystemExit("bench_frame.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        from tools import make_standin_scenes as ms
        ms.generate(detail=2, quiet=True)
    if dist:
        dist.barrier()
    pkg = importlib.import_module("bidirectional-path-tracer_b200")
    W, H = (int(x) for x in args.res.split("x"))
    sf = pkg.SceneFile(args.scene)
    cam = sf.camera_with(W, H, args.spp, args.depth)
    n = int(np.sqrt(float(args.spp)))
    with pkg.SceneImage.build(sf.verts, sf.mat_id, sf.materials) as built:
        pinned = torch.empty(built.nbytes, dtype=torch.uint8, pin_memory=True)
        pinned.numpy()[:] = built.bytes()
    image = pkg.SceneImage.from_bytes(pinned.numpy())
    film_buf = torch.empty((H, W, 3), dtype=torch.float64, pin_memory=True).numpy() if rank == 0 else None

    def sync_all():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
            torch.cuda.synchronize()

    with pkg.Context(local) as ctx:
        # warm-up: a few strata, the NCCL communicator, the workspace allocation
        ctx.upload_image(image)
        ctx.render(cam, seed=args.seed, bvh=pkg.BVH_FAST, strata=(rank, min(n * n, 4 * world), world), clear=True, height=H)
        ptr, nd = ctx.film_device_ptr()
        film_t = torch.as_tensor(pkg.CudaArray(ptr, nd), device=torch.device("cuda", local))
        if dist:
            dist.reduce(film_t, dst=0)
        ctx.stats_reset()
        sync_all()
        t0 = time.perf_counter()
        ctx.upload_image(image)
        ctx.render(cam, seed=args.seed, bvh=pkg.BVH_FAST, strata=pkg.shard_strata(rank, world, args.spp), clear=True, height=H)
        dev_ms = ctx.last_device_ms()
        t_r = time.perf_counter()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        red_ms = 0.0
        if dist:
            ptr, nd = ctx.film_device_ptr()
            film_t = torch.as_tensor(pkg.CudaArray(ptr, nd), device=torch.device("cuda", local))
            ev0.record()
            dist.reduce(film_t, dst=0)
            ev1.record()
            torch.cuda.synchronize()
            red_ms = ev0.elapsed_time(ev1)
        film = ctx.film(out=film_buf) if rank == 0 else None
        sync_all()
        wall = time.perf_counter() - t0
        st = ctx.stats()
        t = torch.tensor([wall, dev_ms + red_ms, dev_ms], dtype=torch.float64, device="cuda")
        tmin = t.clone()
        rays = torch.tensor([st["rays_closest"], st["rays_shadow"], st["samples"]], dtype=torch.float64, device="cuda")
        if dist:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
            dist.all_reduce(rays, op=dist.ReduceOp.SUM)
        if rank == 0:
            samples = W * H * n * n
            assert int(rays[2]) == samples, (int(rays[2]), samples)
            mean = film.reshape(-1, 3).mean(axis=0) / (n * n)
            line = {"workload": "%s %dx%d, %d spp (%dx%d strata), max depth %d, stand-in geometry; one whole frame, strata sharded over %d GPU(s), one NCCL film reduce"
                                % (os.path.relpath(args.scene, ROOT), W, H, n * n, n, n, args.depth, world),
                    "n_gpus": world, "scaling": "strong", "samples": samples, "frame_seconds_wall_max_over_ranks": float(t[0]),
                    "frame_seconds_device_max_over_ranks": float(t[1]) / 1e3, "msamples_per_s": samples / float(t[1]) / 1e3,
                    "msamples_per_s_wall": samples / float(t[0]) / 1e6, "mrays_per_s": float(rays[0] + rays[1]) / float(t[1]) / 1e3,
                    "render_device_ms_slowest_and_fastest_rank": [float(t[2]), float(tmin[2])], "film_reduce_ms": red_ms,
                    "film_bytes": int(film.nbytes), "film_mean_rgb": [float(x) for x in mean],
                    "film_checksum": float(np.nan_to_num(film).sum())}
            print(json.dumps(line), flush=True)
            if args.png:
                pkg.write_png(args.png, ctx.film_rgb8(n * n))
    if dist:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
