"""CPU-side checks of bench.py's measurement contract: the block decompositions of the weak / strong scaling runs and
the reference arm's OpenMP team under a launcher that exports OMP_NUM_THREADS=1 (torchrun does; round 1's reference
arm silently ran single-threaded at N > 1)."""
import json
import os
import subprocess
import sys
import types

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _args(**kw):
    a = dict(gpus=1, workload="poisson7", scaling="weak", level=None, blocks=None)
    a.update(kw)
    return types.SimpleNamespace(**a)


def test_decompositions():
    sys.path.insert(0, ROOT)
    import bench
    d1 = bench.decomposition(_args())
    assert d1["level"] == 9 and d1["blocks"] == (1, 1, 1)  # BASELINE configs[1]
    for n in (2, 4, 8):
        d = bench.decomposition(_args(gpus=n))
        assert d["level"] == 10 and d["blocks"][0] * d["blocks"][1] * d["blocks"][2] == n  # configs[4], weak
        assert d["domain_hi"] == tuple(float(b) for b in d["blocks"])  # the mesh width of one block
        cfg = bench.workload_config(d)
        assert cfg["cells_per_gpu"] == 1024 ** 3 and "configs[4], weak" in cfg["workload"]
    for n in (1, 2, 4, 8):
        d = bench.decomposition(_args(gpus=n, scaling="strong"))
        total = [d["blocks"][k] * d["block_cells"][k] << d["level"] for k in range(3)]
        assert total == [1024, 1024, 1024]
        # the same level hierarchy as the one-block run: 4 cells per dimension on the coarsest level
        assert [d["blocks"][k] * d["block_cells"][k] << d["min_level"] for k in range(3)] == [4, 4, 4]
    d = bench.decomposition(_args(gpus=8, blocks=(2, 2, 2)))
    assert d["blocks"] == (2, 2, 2) and d["level"] == 10


@pytest.mark.slow
def test_reference_arm_threads_under_launcher_env():
    env = dict(os.environ, OMP_NUM_THREADS="1", RANK="0", WORLD_SIZE="2")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                          "--warmup", "1", "--cpu-level", "5"], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2
    cores = len(os.sched_getaffinity(0))
    assert line["cpu_baseline"]["cores"] == cores  # not the launcher's 1
    assert line["cpu_baseline"]["dof"] == (64 - 1) * (32 - 1) * (32 - 1) or line["cpu_baseline"]["dof"] == 31 * 31 * 63
    # rank 1 prints nothing and exits 0
    env["RANK"] = "1"
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                          "--warmup", "1", "--cpu-level", "5"], env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and out.stdout.strip() == ""
