"""Run under torchrun (gloo, CPU): the host-side logic of stratum sharding and the film reduce."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def fake_stratum_film(s, shape):
    """A deterministic stand-in for 'the film of stratum s' (no GPU here): what matters is who adds which s."""
    rng = np.random.default_rng(1000 + s)
    return rng.random(shape)


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    pkg = importlib.import_module("bidirectional-path-tracer_b200")
    shape = (5, 7, 3)
    for spp in (1, 4, 16, 50, 256, 1024):
        n2 = int(np.sqrt(spp)) ** 2
        b, e, st = pkg.shard_strata(rank, world, spp)
        mine = list(range(b, e, st))
        film = np.zeros(shape)
        for s in mine:
            film += fake_stratum_film(s, shape)
        # every stratum belongs to exactly one rank
        owned = [None] * world
        dist.all_gather_object(owned, mine)
        flat = sorted(s for o in owned for s in o)
        assert flat == list(range(n2)), (spp, flat[:8])
        assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1
        t = torch.from_numpy(film)
        pkg.reduce_film(t, dist, dst=0)
        if rank == 0:
            whole = np.zeros(shape)
            for s in range(n2):
                whole += fake_stratum_film(s, shape)
            assert np.allclose(t.numpy(), whole, rtol=1e-13, atol=0), spp
    dist.barrier()
    if rank == 0:
        print("SHARD_OK world=%d" % world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
