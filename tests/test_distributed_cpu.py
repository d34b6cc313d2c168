"""Host-side logic of the multi-GPU path, run on CPU (gloo, world size 2 and 4) -- the GPU data path
itself needs one GPU per rank and is checked by tests/test_multi_gpu.py / tests/mgpu/worker.py."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_decomposition_helpers():
    from exastencils_b200 import distributed as D
    for world, nb in D.BLOCK_GRIDS.items():
        assert nb[0] * nb[1] * nb[2] == world
        seen = set()
        for r in range(world):
            idx = D.block_index(r, nb)
            assert D.block_rank(idx, nb) == r          # x fastest, IR_InitGeneratedDomain.scala:41-47
            seen.add(idx)
            for (d, side), other in D.neighbours(r, nb).items():
                j = list(idx)
                j[d] += 1 if side else -1
                assert D.block_index(other, nb) == tuple(j)
        assert len(seen) == world
    assert D.block_index(5, (2, 2, 2)) == (1, 0, 1)
    assert D.neighbours(0, (1, 1, 1)) == {}


@pytest.mark.parametrize("world", [2, 4])
def test_wiring_over_gloo(world):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    port = 29600 + world
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "mgpu", "cpu_worker.py")]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=240)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    for r in range(world):
        assert "CPU WORKER %d OK" % r in out.stdout
