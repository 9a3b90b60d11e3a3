#!/usr/bin/env python
"""bench.py — the hot path's headline benchmark (BASELINE.json: "BDPT Msamples/s and BVH Mrays/s").

Workload (config.workload): scenes/coffee_machine.yaml at 1920x1080, 256 spp (16x16 strata), max depth 16,
stand-in geometry (the reference ships no OBJ assets).  One STEP = one pass of the whole
wavefront pipeline (generate -> trace -> shade -> connect -> shadow -> film) over STRATA_PER_STEP strata of
every pixel of the frame; consecutive steps render consecutive strata of the 256-spp frame, so no step
repeats another's samples.  With N GPUs each rank renders its own strata (rank g takes stratum s*N+g:
stratum sharding), per-GPU work fixed -> weak scaling; the rank films are summed with ONE NCCL
reduce at the end of the timed region (one reduce per frame).

  python bench.py [--gpus N] [--steps K] [--warmup W]            our CUDA path (libbdpt.so)
  python bench.py --impl reference ...                          the reference's own CPU code (oracle/_ref)

value  = samples of all ranks / device time of the K steps (inputs resident in HBM).
e2e    = same metric through the host-facing call sequence: scene upload from host buffers (BVH build +
         H2D), K steps, film read-back (D2H) — everything a camera::render call pays.  Median of three passes,
         each on a fresh context.
e2e_scene_image = the same with the scene handed over as a prebuilt scene image in pinned host memory
         (bdpt_scene_upload_image: one allocation + one H2D copy inside the timed region; flatten and tree
         builds outside it, as the reference arm's BVH build is).  Reported beside e2e, never instead of it.
bvh_microbench = BASELINE.json config 2 (16,777,216 fixed-seed rays): device-resident, pinned host buffers
         through the C-ABI's chunked copy/trace pipeline, and pageable host buffers.
clocks = SM clock and throttle reasons sampled through NVML every 10 ms during the timed region (rank 0).
"""
import argparse
import ctypes
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

_JSON_OUT = None

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WIDTH, HEIGHT, SPP, DEPTH = 1920, 1080, 256, 16
STRATA_PER_STEP = 4
SCENE = os.path.join(ROOT, "scenes", "coffee_machine.yaml")
MICRO_SCENE = os.path.join(ROOT, "scenes", "glass-of-water.yaml")
MICRO_RAYS = 1 << 24


def ensure_scenes():
    from tools import make_standin_scenes as ms
    ms.generate(detail=2, quiet=True)


def measured_peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons DURING the timed region (B200_PROFILING.md clocks line), read through NVML in
    this process (nvidia_ml_py) every 10 ms; one `nvidia-smi` child per sample only if NVML cannot be loaded.  Under
    torchrun only rank 0 samples: eight ranks each forking nvidia-smi (which enumerates every GPU of the box) stalled
    kernel launches in the driver and cost the 8-GPU run 15 % of its step time."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    BITS = [0x8, 0x40, 0x20, 0x4]   # nvmlClocksEventReason{HwSlowdown, HwThermalSlowdown, SwThermalSlowdown, SwPowerCap}

    def __init__(self, index, enabled=True):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt, self.enabled = index, [], threading.Event(), enabled
        self.nvml, self.handle, self.max_mhz, self.source = None, None, None, "nvidia-smi"
        if not enabled:
            return
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = int(vis.split(",")[index]) if vis and all(x.strip().isdigit() for x in vis.split(",")) else index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml, self.source = pynvml, "nvml"
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n = self.nvml
        mhz = int(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
        try:
            mask = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
        except Exception:
            mask = int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
        self.rows.append((mhz, mask))

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                              "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
        parts = [x.strip() for x in out.strip().split(",")]
        if len(parts) >= 6 and parts[0].isdigit():
            mask = sum(b for k, b in enumerate(self.BITS) if parts[2 + k].lower().startswith("active"))
            self.rows.append((int(parts[0]), mask))
            if parts[1].isdigit():
                self.max_mhz = int(parts[1])

    def start(self):
        if self.enabled:
            super().start()

    def run(self):
        while not self._stop_evt.is_set():
            try:
                self._sample_nvml() if self.nvml else self._sample_smi()
            except Exception:
                pass
            self._stop_evt.wait(0.01 if self.nvml else 0.05)

    def stop(self):
        if not self.enabled:
            return None
        self._stop_evt.set()
        self.join(timeout=6)
        sm = sorted(r[0] for r in self.rows)
        reasons = [n for n, b in zip(self.NAMES, self.BITS) if any(r[1] & b for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": reasons, "samples": len(self.rows), "source": self.source}


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def dump_soup():
    """The workload's triangle soup as an .npz, written by a child process (tools/dump_soup.py): the reference arm's
    own processes then never map libbdpt.so."""
    ensure_scenes()
    out = os.path.join(ROOT, "gpurun_out" if os.path.isdir(os.path.join(ROOT, "gpurun_out")) else ".", "_soup_coffee_machine.npz")
    out = os.path.abspath(out)
    if not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(SCENE):
        subprocess.check_call([sys.executable, os.path.join(ROOT, "tools", "dump_soup.py"), SCENE, out], stdout=subprocess.DEVNULL)
    return out


def cpu_reference(spp, steps, warmup, procs=None, rows=2):
    """The reference CPU renderer on this box's host cores: `procs` independent single-thread processes, each
    rendering `rows` full-width rows of the workload per step at `spp` samples per pixel; every worker's rows are spread
    over the image height and interleaved with the others', so all of them see the same mix of cheap and dear rows.
    The rate is the SUM of the processes' own rates (samples_i / seconds_i): a core that finishes early is not
    counted as idle."""
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "-s"], stdout=subprocess.DEVNULL)
    soup = dump_soup()
    procs = procs or host_cores()
    cmd = [sys.executable, os.path.join(ROOT, "tools", "cpu_ref_worker.py"), soup, str(WIDTH), str(HEIGHT), str(spp), str(DEPTH)]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    ps = [subprocess.Popen(cmd + [str(p), str(procs), str(steps), str(warmup), "1", str(rows)], stdout=subprocess.PIPE,
                           stderr=subprocess.PIPE, text=True, env=env) for p in range(procs)]
    outs, failed = [], []
    for k, p in enumerate(ps):
        out, err = p.communicate()
        lines = [l for l in out.splitlines() if l.startswith("{")]
        if p.returncode == 0 and lines:
            outs.append(json.loads(lines[-1]))
        else:
            failed.append("worker %d rc=%s: %s" % (k, p.returncode, err.strip().splitlines()[-1] if err.strip() else ""))
    if failed:
        raise RuntimeError("CPU reference workers failed: " + "; ".join(failed[:3]))
    samples = sum(o["samples_per_step"] * len(o["step_s"]) for o in outs)
    rate = sum(o["samples_per_step"] * len(o["step_s"]) / sum(o["step_s"]) for o in outs)      # samples/s, summed over the cores
    wall = max(sum(o["step_s"]) for o in outs)
    rays = sum(o["rays"] for o in outs)
    all_samples = sum(o["samples_per_step"] * o["all_steps"] for o in outs)
    return {"value": rate / 1e6, "unit": "Msamples/s", "cores": len(outs), "kind": outs[0]["kind"],
            "sample": "%d independent single-thread processes x %d step(s) x %d interleaved row(s) of %dx%d at %d spp (%dx%d strata), depth %d "
                      "(%d samples timed; per-sample cost does not depend on the stratum count); sum of the processes' rates; "
                      "reference headers compiled -O2, BVH build excluded"
                      % (len(outs), steps, rows, WIDTH, HEIGHT, spp, int(spp ** 0.5), int(spp ** 0.5), DEPTH, samples),
            "ms_per_step": 1e3 * wall / max(1, steps), "closest_rays_per_sample": rays / max(1, all_samples),
            "node_visits_per_ray": sum(o["node_visits"] for o in outs) / max(1, rays),
            "tri_tests_per_ray": sum(o["tri_tests"] for o in outs) / max(1, rays),
            "slowest_over_fastest_worker": max(sum(o["step_s"]) for o in outs) / max(1e-9, min(sum(o["step_s"]) for o in outs))}


def cpu_reference_variant(spp, threads, rows):
    """One process of the reference through tools/cpu_ref_worker.py: threads > 1 = its own thread-pool shape."""
    cmd = [sys.executable, os.path.join(ROOT, "tools", "cpu_ref_worker.py"), dump_soup(), str(WIDTH), str(HEIGHT), str(spp), str(DEPTH),
           "0", "1", "1", "0", str(threads), str(rows)]
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, OMP_NUM_THREADS="1")).stdout
        o = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
        return {"value": o["samples_per_step"] / sum(o["step_s"]) / 1e6, "unit": "Msamples/s", "threads": threads,
                "sample": "%d row(s) of %dx%d at %d spp, depth %d (%d samples)" % (rows, WIDTH, HEIGHT, spp, DEPTH, o["samples_per_step"])}
    except Exception as e:   # reporting only: never fail the arm over it
        return {"unavailable": str(e)[:200]}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # a row at 16 spp (4x4 strata) costs ~1 s per process; two interleaved rows per step keep a step bounded however many
    # steps are asked for.  The workload's 256 spp differ only in the number of strata: the cost per SAMPLE is the same.
    r = cpu_reference(16, args.steps, args.warmup, rows=2)
    cfg = dict(workload_config(), reference_sample="each step: 2 interleaved rows per process at 16 spp (4x4 strata of the workload's 16x16)")
    line = {"impl": "reference", "metric": "BDPT Msamples/s", "value": r["value"], "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": cfg,
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "reference_counters": {k: r[k] for k in ("closest_rays_per_sample", "node_visits_per_ray", "tri_tests_per_ray", "slowest_over_fastest_worker")}}
    # the other two ways of running the reference on these cores, reported beside the headline (iii)
    if not args.skip_cpu:
        cores = host_cores()
        line["variants"] = {
            "independent_processes": "the line's value (one single-thread process per core, interleaved rows)",
            "single_thread": cpu_reference_variant(16, 1, 4),
            "as_shipped_thread_pool": dict(cpu_reference_variant(16, cores + 1, 8),
                                           note="hardware_concurrency()+1 threads sharing one unlocked mt19937 and the stats atomics "
                                                "(camera.h:60, main.h:28-32, stats.h:8-16): what a user of the reference gets; racy, so "
                                                "its image differs from run to run")}
    emit(line)
    return 0


def workload_config():
    return {"workload": "scenes/coffee_machine.yaml %dx%d, %d spp (16x16 strata), max depth %d, stand-in geometry "
                        "(reference ships no OBJ assets); step = %d strata of every pixel"
                        % (WIDTH, HEIGHT, SPP, DEPTH, STRATA_PER_STEP),
            "samples_per_step_per_gpu": WIDTH * HEIGHT * STRATA_PER_STEP,
            "l2": "working set per step (wavefront state of 4 workers x 2^21 sample slots, ~46 GB of path/vertex records and queues, streamed every iteration) exceeds the 126 MB L2; "
                  "no flush needed; the 50 MB scene itself is L2-resident by design",
            "parallelism": "stratum sharding, one NCCL film reduce per frame"}


class CudaArray:
    """__cuda_array_interface__ view of a device pointer owned by libbdpt (the film), so torch.distributed
    can reduce it in place over NCCL."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 3}


def run_ours(args):
    import torch
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and "BDPT_REF_FORKS" not in os.environ:
        # several ranks build their reference-topology trees side by side: cores / world subtrees in flight per rank (the
        # library's default of 16 assumes the box to itself)
        per_rank = max(2, (os.cpu_count() or 16) // world)
        os.environ["BDPT_REF_FORKS"] = str(max(1, min(4, per_rank.bit_length() - 1)))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        ensure_scenes()
    if dist:
        dist.barrier()
    pkg = importlib.import_module("bidirectional-path-tracer_b200")
    sf = pkg.SceneFile(SCENE)
    cam = sf.camera_with(WIDTH, HEIGHT, SPP, DEPTH)
    bvh = pkg.BVH_REFERENCE if args.bvh == "ref" else pkg.BVH_FAST
    ctx = pkg.Context(local)
    K, Wm = args.steps, args.warmup
    n2 = 256

    def strata_of(step):          # this rank's strata for a step; wraps after the 256-spp frame is complete
        base = ((step * world + rank) * STRATA_PER_STEP) % n2
        return (base, base + STRATA_PER_STEP, 1)

    def sync_all():
        ctx.sync()
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
            torch.cuda.synchronize()

    # ---------------- device-resident measurement
    ctx.upload(sf.verts, sf.mat_id, sf.materials, gpu_tree=args.gpu_tree)
    ctx.film_clear(WIDTH, HEIGHT)
    if dist:
        pkg.join_comm(ctx, dist)     # the film reduce is the library's own (bdpt_film_reduce_comm over NCCL); torch only carries the id
    for s in range(Wm):
        ctx.render(cam, seed=args.seed, bvh=bvh, strata=strata_of(s), clear=False, height=HEIGHT, batch_samples=args.batch, streams=args.streams)
    if dist:
        ctx.film_reduce(0)           # warm NCCL up too
    ctx.film_clear(WIDTH, HEIGHT)
    ctx.stats_reset()
    sampler = ClockSampler(local, enabled=(rank == 0))
    sync_all()
    sampler.start()
    t0 = time.perf_counter()
    dev_ms = 0.0
    for s in range(K):
        ctx.render(cam, seed=args.seed, bvh=bvh, strata=strata_of(Wm + s), clear=False, height=HEIGHT, batch_samples=args.batch, streams=args.streams)
        dev_ms += ctx.last_device_ms()
    red_ms = 0.0
    if dist:
        red_ms = ctx.film_reduce(0)  # the one film reduce per frame: device time of ncclReduce on the library's stream
    sync_all()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop()
    st = ctx.stats()
    kt = ctx.kernel_times()
    step_ms = torch.tensor([dev_ms + red_ms, wall_ms], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(step_ms, op=dist.ReduceOp.MAX)
    total_ms, wall_ms = float(step_ms[0]), float(step_ms[1])
    samples = WIDTH * HEIGHT * STRATA_PER_STEP * K * world
    value = samples / (total_ms * 1e3)

    # ---------------- end-to-end through host buffers: upload (BVH build + H2D) + K steps + film D2H
    # Three passes, each with a fresh context; the MEDIAN pass is reported (all three are kept in the line): a
    # sporadic 20-80 ms stall inside a driver call (cudaMalloc of the 46 GB workspace, the first H2D of a new
    # context) otherwise decides a 250 ms measurement.
    film_buf = torch.empty((HEIGHT, WIDTH, 3), dtype=torch.float64, pin_memory=True).numpy()   # the caller's frame buffer, pinned

    shm = "/dev/shm/bdpt_bench_%s.bdptimg" % os.environ.get("MASTER_PORT", "0")

    def upload_cold(c):
        """From the flattened soup in host buffers: bdpt_scene_upload on every rank (flatten + the reference-topology tree on
        the host, the fast tree on the rank's own device, H2D).  With several ranks the host builds share the box's cores, so
        each rank's reference build forks into cores / world subtrees (BDPT_REF_FORKS, set in main before the first upload).
        BENCH_E2E_SHARED_IMAGE=1 gives round 2's first scheme instead: rank 0 builds ONE scene image (both trees on the host)
        into shared memory and every rank uploads that block -- what `bdpt --gpus N` does inside one process."""
        if not dist or not os.environ.get("BENCH_E2E_SHARED_IMAGE"):
            return c.upload(sf.verts, sf.mat_id, sf.materials, gpu_tree=args.gpu_tree)
        if rank == 0:
            with pkg.SceneImage.build(sf.verts, sf.mat_id, sf.materials) as img:
                img.save(shm)
        dist.barrier()
        with pkg.SceneImage.map(shm) as img:
            c.upload_image(img)

    def e2e_passes(upload):
        passes = []
        film_host = None
        for _ in range(int(os.environ.get("BENCH_E2E_PASSES", "3"))):
            ctx2 = pkg.Context(local)
            if dist:
                ctx2.comm_attach()   # the process joined the communicator once, above: a frame does not pay a rendezvous
            sync_all()
            t0 = time.perf_counter()
            upload(ctx2)
            t_up = time.perf_counter()
            ctx2.film_clear(WIDTH, HEIGHT)
            per_call = []
            for s in range(K):
                tc = time.perf_counter()
                ctx2.render(cam, seed=args.seed, bvh=bvh, strata=strata_of(Wm + s), clear=False, height=HEIGHT, batch_samples=args.batch, streams=args.streams)
                per_call.append(round((time.perf_counter() - tc) * 1e3, 2))
            t_rd = time.perf_counter()
            if dist:
                ctx2.film_reduce(0)
            film_host = ctx2.film(out=film_buf)
            sync_all()
            t_end = time.perf_counter()
            e2e_s = torch.tensor([t_end - t0], dtype=torch.float64, device="cuda")
            if dist:
                dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
            passes.append({"seconds_max_over_ranks": float(e2e_s[0]), "scene_upload_ms": (t_up - t0) * 1e3, "render_ms": (t_rd - t_up) * 1e3,
                           "film_reduce_and_readback_ms": (t_end - t_rd) * 1e3, "render_calls_ms": per_call})
            ctx2.close()
        return passes, sorted(passes, key=lambda p: p["seconds_max_over_ranks"])[len(passes) // 2], film_host

    passes, e2e_parts, film_host = e2e_passes(upload_cold)
    if dist and rank == 0 and os.path.exists(shm):
        os.unlink(shm)
    e2e_value = samples / e2e_parts["seconds_max_over_ranks"] / 1e6
    film_checksum = float(np.nan_to_num(film_host).sum())

    # The same with the scene handed over as a prebuilt scene image (bdpt_scene_image_build once, outside the timed
    # region, kept in pinned host memory; bdpt_scene_upload_image = one allocation + one H2D copy inside it): what a
    # caller that renders a scene more than once, or keeps the image file beside the scene, pays per frame.  The
    # reference arm's timed region excludes its BVH build too.
    t_b = time.perf_counter()
    with pkg.SceneImage.build(sf.verts, sf.mat_id, sf.materials) as built:
        image_build_ms = (time.perf_counter() - t_b) * 1e3
        pinned = torch.empty(built.nbytes, dtype=torch.uint8, pin_memory=True)
        pinned.numpy()[:] = built.bytes()
    image = pkg.SceneImage.from_bytes(pinned.numpy())
    image_info = image.info(verify=True)
    img_passes, img_parts, film_img = e2e_passes(lambda c: c.upload_image(image))
    img_value = samples / img_parts["seconds_max_over_ranks"] / 1e6
    img_checksum = float(np.nan_to_num(film_img).sum())
    h2d = sf.verts.nbytes + sf.mat_id.nbytes
    d2h = film_host.nbytes

    frame = None
    if not args.skip_frame:
        frame = frame_leg(args, pkg, ctx, sf, bvh, dist, rank, world, sync_all)
    extra = {}
    if rank == 0:
        extra = rank0_extras(args, pkg, ctx, sf, cam, bvh, st, kt, K)
        if frame:
            extra["config4_frame"] = frame
    ctx.close()
    if rank == 0:
        line = {"metric": "BDPT Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": K, "warmup": Wm,
                "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic", "config": dict(workload_config(), bvh=args.bvh, fast_tree="host, binned SAH" if args.gpu_tree is False else "device, binned SAH (library default)"),
                "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": h2d // K, "d2h_bytes_per_step": d2h // K,
                        "rank0_ms": e2e_parts, "all_passes_msamples_per_s": [samples / p["seconds_max_over_ranks"] / 1e6 for p in passes],
                        "all_passes_rank0_ms": passes,
                        "note": "scene upload from host buffers (flatten + both host BVH builds + H2D; several ranks: rank 0 builds one "
                                "scene image in shared memory, all upload it) + K steps + NCCL film reduce + film read-back into a pinned host "
                                "frame buffer; median of three passes, each on a fresh context"},
                "gpu_launches": st["kernel_launches"], "clocks": clocks, "wall_ms_per_step": wall_ms / K,
                "film_reduce_ms": red_ms,
                "mrays_per_s": (st["rays_closest"] + st["rays_shadow"]) * world / (total_ms * 1e3),
                "rays_per_sample": {"closest": st["rays_closest"] / st["samples"], "shadow": st["rays_shadow"] / st["samples"]},
                "e2e_scene_image": {"value": img_value, "unit": "Msamples/s", "h2d_bytes_per_step": image_info["device_bytes"] // K,
                                    "d2h_bytes_per_step": d2h // K, "rank0_ms": img_parts,
                                    "all_passes_msamples_per_s": [samples / p["seconds_max_over_ranks"] / 1e6 for p in img_passes],
                                    "image_bytes": image.nbytes, "image_build_ms_outside_timed_region": image_build_ms,
                                    "film_checksum": img_checksum,
                                    "note": "as e2e, but the scene arrives as a prebuilt scene image in pinned host memory "
                                            "(bdpt_scene_upload_image: one cudaMalloc + one H2D copy); flatten and both BVH builds are "
                                            "outside the timed region, as the reference arm's BVH build is"},
                "film_checksum": film_checksum}
        line.update(extra)
        emit(line)
    if dist:
        dist.barrier()
        dist.destroy_process_group()
    return 0


FRAME_W, FRAME_H, FRAME_SPP = 3840, 2160, 1024


def frame_leg(args, pkg, ctx, sf, bvh, dist, rank, world, sync_all):
    """BASELINE.json config 4: ONE whole coffee_machine frame at 3840x2160, 1024 spp (32x32 strata), depth 16, its strata
    sharded over the ranks (rank g renders strata g, g+N, ...), the NCCL film reduce inside the timed region, then the
    read-back of the summed film on rank 0.  Total work is fixed -> the per-N values of this leg are STRONG scaling."""
    import torch
    cam4 = sf.camera_with(FRAME_W, FRAME_H, FRAME_SPP, DEPTH)
    n2 = int(FRAME_SPP ** 0.5) ** 2
    ctx.render(cam4, seed=args.seed, bvh=bvh, strata=(rank, min(n2, world), world), clear=True, height=FRAME_H)   # warm-up: one stratum (sizes the film)
    if dist:
        ctx.film_reduce(0)
    host = torch.empty((FRAME_H, FRAME_W, 3), dtype=torch.float64, pin_memory=True).numpy() if rank == 0 else None
    sync_all()
    t0 = time.perf_counter()
    ctx.render(cam4, seed=args.seed, bvh=bvh, strata=pkg.shard_strata(rank, world, FRAME_SPP), clear=True, height=FRAME_H)
    render_ms = ctx.last_device_ms()
    red_ms = ctx.film_reduce(0) if dist else 0.0
    t_dev = time.perf_counter()
    if rank == 0:
        ctx.film(out=host)
    sync_all()
    t1 = time.perf_counter()
    st = ctx.stats()
    v = torch.tensor([render_ms + red_ms, t_dev - t0, t1 - t0, render_ms], dtype=torch.float64, device="cuda")
    lo = v.clone()
    if dist:
        dist.all_reduce(v, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    samples = FRAME_W * FRAME_H * n2
    return {"workload": "scenes/coffee_machine.yaml %dx%d, %d spp, max depth %d: one whole frame, strata sharded over %d GPU(s), film reduce inside"
                        % (FRAME_W, FRAME_H, FRAME_SPP, DEPTH, world),
            "scaling": "strong", "samples": samples, "frame_s": float(v[0]) / 1e3, "msamples_per_s": samples / (float(v[0]) * 1e3),
            "film_reduce_ms": red_ms, "film_bytes": FRAME_W * FRAME_H * 24,
            "render_ms_slowest_rank": float(v[3]), "render_ms_fastest_rank": float(lo[3]),
            "wall_s_render_and_reduce": float(v[1]), "e2e_frame_s_with_readback": float(v[2]),
            "e2e_msamples_per_s": samples / float(v[2]) / 1e6,
            "film_checksum": float(np.nan_to_num(host).sum()) if rank == 0 else None,
            "timing": "device: CUDA events around the render + the reduce, max over ranks; e2e: wall clock, barrier on both sides"}


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel`, from the committed ncu --set full
    summary of this round (profiles/): a number measured under the profiler, quoted, not re-measured here."""
    path = os.path.join(ROOT, "profiles", "r2_ncu_full_%s.csv" % kernel)   # this round's capture of the build being benchmarked
    try:
        vals = {}
        for line in open(path):
            p = line.strip().split(",")
            if len(p) >= 3 and p[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                vals[p[0]] = float(p[1]) * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[p[2]]
        if len(vals) == 2:
            return sum(vals.values()), "bytes of the launch captured in profiles/%s" % os.path.basename(path)
    except Exception:
        pass
    return None, "no ncu capture committed for this kernel"


def rank0_extras(args, pkg, ctx, sf, cam, bvh, st, kt, K):
    """roofline of the dominant kernel class, the 16M-ray BVH microbench (config 2) and the CPU baseline."""
    out = {}
    peak, peak_src = measured_peaks()
    # algorithmic bytes per ray: SURVEY 8(d)  A_ray = 32 (ray in) + 16 (hit out) + 32 x nodes + 48 x triangles,
    # nodes / triangles counted by the SAME kernels in a stats pass over one step's strata (untimed)
    ctx.stats_reset()
    ctx.render(cam, seed=args.seed, bvh=bvh, strata=(0, STRATA_PER_STEP, 1), clear=True, height=HEIGHT, stats=True)
    s2 = ctx.stats()
    per = {}
    for kind, rays, nodes, tris in (("trace", "rays_closest", "nodes_closest", "tris_closest"),
                                    ("shadow", "rays_shadow", "nodes_shadow", "tris_shadow")):
        r = max(1, s2[rays])
        per[kind] = {"nodes_per_ray": s2[nodes] / r, "tris_per_ray": s2[tris] / r,
                     "bytes_per_ray": 48.0 + 32.0 * s2[nodes] / r + 48.0 * s2[tris] / r}
    # per-kernel durations: CUDA-event pairs around every launch, in a pass with ONE wavefront in flight
    # (streams=1) so that a launch's duration is its own and not its share of a GPU it co-runs on; the
    # timed region above keeps several wavefronts in flight and is faster than the sum of these
    ctx.stats_reset()
    for s_ in range(2):
        ctx.render(cam, seed=args.seed, bvh=bvh, strata=((s_ + 1) * STRATA_PER_STEP, (s_ + 2) * STRATA_PER_STEP, 1),
                   clear=False, height=HEIGHT, streams=1, serial=True)
    st, kt = ctx.stats(), ctx.kernel_times()
    out["serialized_pass"] = {"streams": 1, "steps": 2, "ms_per_step": sum(v[0] for v in kt.values()) / 2}
    dom = "trace" if kt["trace"][0] >= kt["shadow"][0] else "shadow"
    rays_dom = st["rays_closest"] if dom == "trace" else st["rays_shadow"]
    ms_dom, launches_dom = kt[dom]
    achieved = rays_dom * per[dom]["bytes_per_ray"] / (ms_dom * 1e-3) / 1e9 if ms_dom > 0 else 0.0
    total_kernel_ms = sum(v[0] for v in kt.values())
    fast = bvh == pkg.BVH_FAST
    kname = ("k_trace_fast" if fast else "k_trace_paths") if dom == "trace" else ("k_shadow_fast" if fast else "k_shadow")
    traffic, traffic_note = ncu_traffic(kname)
    try:   # the named scenes are L2-resident by design: the L2 read peak (48 MiB buffer, 16-byte loads) beside the HBM figure
        l2_peak = ctx.bench_read(48 << 20, reps=40)
        hbm_read = ctx.bench_read(4 << 30, reps=4)
    except Exception:
        l2_peak, hbm_read = None, None
    out["roofline"] = {"bound": "hbm", "kernel": kname, "achieved": achieved,
                       "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_note": traffic_note,
                       "peak_source": peak_src,
                       "l2_peak": l2_peak, "l2_frac": (achieved / l2_peak) if l2_peak else None, "hbm_read_probe": hbm_read,
                       "l2_note": "l2_peak: bdpt_bench_read over a 48 MiB resident buffer, measured in this run; the algorithmic bytes of "
                                  "this scene are served by L1/L2 (the 60 MB scene is L2-resident), so l2_frac is the bound that applies and "
                                  "frac (of HBM) a figure of merit; hbm_read_probe: the same probe over 4 GiB",
                       "avg_launch_ms": ms_dom / max(1, launches_dom), "launches": launches_dom,
                       "algorithmic_bytes_per_ray": per[dom]["bytes_per_ray"],
                       "algorithmic_bytes_per_launch": rays_dom * per[dom]["bytes_per_ray"] / max(1, launches_dom),
                       "mrays_per_s_in_kernel": rays_dom / (ms_dom * 1e3) if ms_dom > 0 else 0.0,
                       "share_of_step": ms_dom / total_kernel_ms if total_kernel_ms else None,
                       "timing": "CUDA events per launch on the launching stream, streams=1 pass"}
    out["kernel_ms"] = {k: {"ms": v[0], "launches": v[1]} for k, v in kt.items()}
    per["max_work_single_ray"] = {"closest": s2["max_work_closest"], "shadow": s2["max_work_shadow"],
                                  "note": "nodes + triangles touched by the most expensive ray of the stats pass"}
    out["traversal_work"] = per
    if not args.skip_micro:
        out["bvh_microbench"] = bvh_microbench(args, pkg, bvh)
    if not args.skip_configs:
        try:
            out.update(config_legs(args, pkg, bvh, peak))
        except Exception as e:   # the headline stands on its own
            out["config_legs_error"] = str(e)[:300]
    if not args.skip_cpu:
        try:
            r = cpu_reference(64, 1, 0)
            out["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        except Exception as e:  # the GPU numbers stand on their own
            out["cpu_baseline"] = {"value": None, "unit": "Msamples/s", "cores": 0, "kind": "reference", "sample": "failed: %s" % e}
    return out


def config_legs(args, pkg, bvh, peak):
    """BASELINE.json configs 1 and 5 as legs of the line (rank 0, after the headline measurement).
    config 1: glass-of-water 512x512, 16 spp, the YAML's own depth (80): one whole frame, device time and end to end.
    config 5: the 10,000,008-triangle sphere grid, 1920x1080, 64 spp, depth 32 -- the one config whose trees do NOT fit
    the L2, i.e. where the HBM roofline really binds: steps of 2 strata, the trace kernel's algorithmic bytes / time."""
    import torch
    from tools import testscenes as ts
    out = {}
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    # ---- config 1
    sf = pkg.SceneFile(MICRO_SCENE)
    cam1 = sf.camera_with(512, 512, 16, None)
    host = torch.empty((512, 512, 3), dtype=torch.float64, pin_memory=True).numpy()
    with pkg.Context(dev) as c:
        c.upload(sf.verts, sf.mat_id, sf.materials)
        c.render(cam1, seed=args.seed, bvh=bvh, height=512)            # warm-up frame
        c.stats_reset()
        ms = []
        for k in range(3):
            c.render(cam1, seed=args.seed + k, bvh=bvh, height=512)
            ms.append(c.last_device_ms())
        st = c.stats()
    t0 = time.perf_counter()
    with pkg.Context(dev) as c:
        c.upload(sf.verts, sf.mat_id, sf.materials)
        c.render(cam1, seed=args.seed, bvh=bvh, height=512)
        c.film(out=host)
    e2e_s = time.perf_counter() - t0
    n1 = 512 * 512 * 16
    out["config1_glass_of_water"] = {
        "workload": "scenes/glass-of-water.yaml 512x512, 16 spp (4x4 strata), max depth %d (the YAML's), stand-in geometry, %d triangles" % (cam1.max_depth, len(sf.verts)),
        "msamples_per_s": n1 / min(ms) / 1e3, "frame_ms": min(ms), "frame_ms_all": ms,
        "e2e_msamples_per_s": n1 / e2e_s / 1e6, "e2e_frame_s": e2e_s,
        "e2e_note": "fresh context: scene upload from host buffers (both BVH builds) + one frame + film read-back; dominated by the ~0.1 s of "
                    "context and workspace set-up at this frame size",
        "rays_per_sample": {"closest": st["rays_closest"] / max(1, st["samples"]), "shadow": st["rays_shadow"] / max(1, st["samples"])},
        "film_checksum": float(np.nan_to_num(host).sum())}
    # ---- config 5
    v, m, mats, camkw = ts.sphere_grid(50)
    W5, H5, SPP5, D5, S5 = 1920, 1080, 64, 32, 2
    cam5 = pkg.make_camera(W5, H5, SPP5, D5, **camkw)
    with pkg.Context(dev) as c:
        t0 = time.perf_counter()
        c.upload(v, m, mats, gpu_tree=args.gpu_tree)
        up_s = time.perf_counter() - t0
        c.film_clear(W5, H5)
        for s_ in range(2):
            c.render(cam5, seed=1, bvh=bvh, strata=(s_ * S5, (s_ + 1) * S5, 1), clear=False, height=H5)
        c.stats_reset()
        ms, steps = 0.0, 4
        for s_ in range(steps):
            c.render(cam5, seed=1, bvh=bvh, strata=((2 + s_) * S5, (3 + s_) * S5, 1), clear=False, height=H5)
            ms += c.last_device_ms()
        st = c.stats()
        c.stats_reset()
        c.render(cam5, seed=1, bvh=bvh, strata=(0, 1, 1), clear=True, height=H5, stats=True)
        s2 = c.stats()
        c.stats_reset()
        c.render(cam5, seed=1, bvh=bvh, strata=(20, 22, 1), clear=True, height=H5, streams=1, serial=True)
        st1, kt = c.stats(), c.kernel_times()
        nodes, fastn = c.bvh_info(pkg.BVH_REFERENCE), c.bvh_info(pkg.BVH_FAST)
    rays = max(1, s2["rays_closest"])
    bpr = 48.0 + 32.0 * s2["nodes_closest"] / rays + 48.0 * s2["tris_closest"] / rays
    tms, tl = kt["trace"]
    ach = st1["rays_closest"] * bpr / (tms * 1e-3) / 1e9 if tms > 0 else 0.0
    samples = W5 * H5 * S5 * steps
    out["config5_ten_million_triangles"] = {
        "workload": "synthetic sphere grid 50x50, %d triangles, %dx%d, %d spp, max depth %d, mixed dielectric / fuzzy metal / lambertian; step = %d strata"
                    % (len(v), W5, H5, SPP5, D5, S5),
        "msamples_per_s": samples / ms / 1e3, "ms_per_step": ms / steps,
        "mrays_per_s": (st["rays_closest"] + st["rays_shadow"]) / ms / 1e3,
        "rays_per_sample": {"closest": st["rays_closest"] / max(1, st["samples"]), "shadow": st["rays_shadow"] / max(1, st["samples"])},
        "bvh_nodes": {"reference_topology": nodes, "fast": fastn}, "scene_upload_s": up_s,
        "roofline": {"bound": "hbm", "kernel": "k_trace_fast", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                     "traffic": ncu_traffic("k_trace_fast_10m_triangles")[0],
                     "algorithmic_bytes_per_ray": bpr, "nodes_per_ray": s2["nodes_closest"] / rays, "tris_per_ray": s2["tris_closest"] / rays,
                     "avg_launch_ms": tms / max(1, tl), "launches": tl, "mrays_per_s_in_kernel": st1["rays_closest"] / (tms * 1e3) if tms > 0 else 0.0,
                     "timing": "CUDA events per launch on the launching stream, streams=1 pass over one step",
                     "note": "the trees (~0.4 GB of nodes, ~0.8 GB of fp64 triangles) do not fit the 126 MB L2: the one config bound by HBM latency/bandwidth"}}
    return out


def bvh_microbench(args, pkg, bvh):
    """BASELINE.json config 2: 16,777,216 fixed-seed rays (camera / incoherent / shadow-segment thirds, fp32-
    representable) against the glass-of-water BVH. Device-resident Mrays/s and host-buffer (e2e) Mrays/s."""
    from tools import testscenes as ts
    sf = pkg.SceneFile(MICRO_SCENE)
    cam = {"lookfrom": tuple(sf.camera.lookfrom), "lookat": tuple(sf.camera.lookat), "vfov": sf.camera.vfov}
    rays = ts.ray_batch(sf.verts, MICRO_RAYS, seed=0x5EED0001, cam=cam, width=512, height=512)
    n = len(rays)
    with pkg.Context(int(os.environ.get("LOCAL_RANK", "0"))) as c:
        c.upload(sf.verts, sf.mat_id, sf.materials, gpu_tree=args.gpu_tree)
        res = {"rays": n, "triangles": len(sf.verts), "scene": "glass-of-water (stand-in geometry)"}
        import torch
        # device-resident: rays and hits live in HBM (torch owns the buffers), nothing is copied in the timed call
        d_rays = torch.from_numpy(np.ascontiguousarray(rays, dtype=np.float64)).cuda()
        d_hits = torch.empty(n * pkg.HIT_DTYPE.itemsize, dtype=torch.uint8, device="cuda")
        # host buffers: the caller's rays and hit records in PINNED memory, through the C-ABI's chunked copy/trace pipeline
        h_rays = torch.empty((n, 8), dtype=torch.float64, pin_memory=True)
        h_rays.numpy()[:] = rays
        h_hits = torch.empty(n * pkg.HIT_DTYPE.itemsize, dtype=torch.uint8, pin_memory=True).numpy().view(pkg.HIT_DTYPE)
        torch.cuda.synchronize()
        for name, flavour in (("fast", pkg.BVH_FAST), ("reference_topology", pkg.BVH_REFERENCE)):
            best_ms, best_wall, best_pageable = 1e30, 1e30, 1e30
            for rep in range(3):
                c.trace_device(d_rays.data_ptr(), n, d_hits.data_ptr(), bvh=flavour)
                best_ms = min(best_ms, c.last_device_ms())
            c.trace(h_rays.numpy()[: 1 << 21], bvh=flavour, out=h_hits[: 1 << 21])   # warm the pinned path (staging, first touch)
            walls = []
            for rep in range(5):
                t0 = time.perf_counter()
                h = c.trace(h_rays.numpy(), bvh=flavour, out=h_hits)
                walls.append(time.perf_counter() - t0)
            best_wall = min(walls)
            for rep in range(2):
                t0 = time.perf_counter()
                hp = c.trace(rays, bvh=flavour)                 # pageable numpy arrays, output allocated by the call
                best_pageable = min(best_pageable, time.perf_counter() - t0)
            dev_prim = d_hits.cpu().numpy().view(pkg.HIT_DTYPE)["prim"]
            assert np.array_equal(dev_prim, h["prim"]) and np.array_equal(hp["prim"], h["prim"])
            c.stats_reset()
            c.trace(rays[: 1 << 20], bvh=flavour, stats=True)
            s = c.stats()
            res[name] = {"mrays_per_s": n / best_ms / 1e3, "e2e_mrays_per_s": n / best_wall / 1e6,
                         "e2e_mrays_per_s_median": n / sorted(walls)[len(walls) // 2] / 1e6,
                         "e2e_pageable_mrays_per_s": n / best_pageable / 1e6, "device_ms": best_ms,
                         "nodes_per_ray": s["node_visits"] / (1 << 20), "tris_per_ray": s["tri_tests"] / (1 << 20),
                         "hit_fraction": float((h["prim"] != 0xFFFFFFFF).mean()),
                         "prim_checksum": int(h["prim"].astype(np.uint64).sum())}
        res["e2e_note"] = ("e2e: rays and hit records in pinned host memory, bdpt_trace copies and traces in 2^20-ray chunks on two streams, best and median of 5 after one warm-up call; "
                           "e2e_pageable: plain numpy arrays, output allocated inside the call")
        res["prim_ids_equal_between_flavours"] = res["fast"]["prim_checksum"] == res["reference_topology"]["prim_checksum"]
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bvh", default="fast", choices=["fast", "ref"],
                    help="traversal flavour; both return the reference's hits (tests/test_trace_gpu.py)")
    ap.add_argument("--seed", type=int, default=20261104)
    ap.add_argument("--batch", type=int, default=0, help="sample slots per wavefront (0 = library default)")
    ap.add_argument("--streams", type=int, default=0, help="wavefronts in flight (0 = library default)")
    ap.add_argument("--host-tree", dest="gpu_tree", action="store_const", const=False, default=None, help="build the fast tree with the host's binned-SAH builder instead of on the device (the library's default)")
    ap.add_argument("--gpu-tree", dest="gpu_tree", action="store_const", const=True, help="build the fast tree on the device (the default)")
    ap.add_argument("--skip-micro", action="store_true")
    ap.add_argument("--skip-frame", action="store_true", help="skip the whole-frame leg (BASELINE.json config 4: 23 s on one GPU)")
    ap.add_argument("--skip-configs", action="store_true", help="skip the config 1 / config 5 legs")
    ap.add_argument("--skip-cpu", action="store_true")
    args = ap.parse_args()
    args.steps = max(1, args.steps)
    args.warmup = max(3, args.warmup) if args.impl == "ours" else max(0, args.warmup)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus != world:
        if world == 1 and args.gpus > 1:   # launched without torchrun: re-exec under it
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(args.gpus),
                   "--master-addr", "127.0.0.1", "--master-port", "29531", os.path.abspath(__file__)] + sys.argv[1:]
            return subprocess.call(cmd)
    # the scene loader prints "Triangles: N" on stdout like scene_loader.h:521; keep stdout for the ONE JSON line
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
