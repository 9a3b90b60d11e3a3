"""Multi-GPU path, host side: stratum sharding + one film reduce, on gloo with 2 and 3 CPU ranks.
The device side of the same path is tests/test_render_gpu.py::test_render_is_independent_of_batching_and_sharding."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world,port", [(2, 29611), (3, 29612)])
def test_strata_partition_and_film_reduce_gloo(world, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "_shard_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=dict(os.environ, OMP_NUM_THREADS="1"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "SHARD_OK world=%d" % world in r.stdout


def test_shard_strata_values(bdpt):
    assert bdpt.shard_strata(0, 1, 256) == (0, 256, 1)
    assert bdpt.shard_strata(3, 8, 1024) == (3, 1024, 8)
    assert bdpt.shard_strata(1, 2, 50) == (1, 49, 2)      # sqrt_spp = 7 (camera.h:164)
    with pytest.raises(ValueError):
        bdpt.shard_strata(2, 2, 16)
