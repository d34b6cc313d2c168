"""Multi-GPU parity: needs >= 2 GPUs (skipped on a single-GPU box).  One rank per GPU runs
tests/mgpu/worker.py: the Benchmark/Poisson3D program on a 2- or 8-block domain against the oracle's
run on the global grid -- fields bit-exact, residual norms within 1e-13."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("world,blocks,level", [(2, "2,1,1", 7), (2, "1,1,2", 7), (4, "1,2,2", 7), (4, "2,2,1", 7), (8, "2,2,2", 6),
                                                (8, "1,2,4", 6)])
def test_blocks_match_global_oracle(world, blocks, level):
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29700 + world), os.path.join(ROOT, "tests", "mgpu", "worker.py"),
           "--level", str(level), "--cycles", "3", "--blocks", blocks]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "MGPU PARITY OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


@pytest.mark.parametrize("world,blocks,level", [(2, "1,1,2", 7), (8, "2,2,2", 7)])
def test_stencil_field_blocks(world, blocks, level):
    """BASELINE config 3's program (27-point stencil field assembled on the device, V(2,2) damped Jacobi with fused
    pairs of sweeps, coefficients of the agglomerated levels re-created on rank 0) on several blocks: bit-identical to
    the same program on one block covering the global grid, and within 1e-9 of the oracle (host exp)."""
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29720 + world), os.path.join(ROOT, "tests", "mgpu", "worker.py"),
           "--level", str(level), "--cycles", "3", "--blocks", blocks, "--op", "varcoeff27", "--smoother", "jacobi"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "MGPU PARITY OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


@pytest.mark.parametrize("world,name,blocks", [(2, "2D_Basic", "2,1"), (2, "2D_Neumann", "1,2"), (2, "3D_Neumann", "1,1,2"),
                                               (2, "3D_Basic", "2,1,1"), (4, "3D_Neumann", "2,2,1"), (8, "3D_Basic", "2,2,2")])
def test_cell_fields_on_several_blocks(world, name, blocks):
    """Cell-localised fields on several blocks: the Testing/CellBased programs split over 2 / 4 / 8 GPUs
    print the reference's golden stdout and give the oracle's global Solution (tests/mgpu/cell_worker.py)."""
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29760 + world), os.path.join(ROOT, "tests", "mgpu", "cell_worker.py"),
           "--name", name, "--blocks", blocks]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "CELL MGPU OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


@pytest.mark.parametrize("variant", ["default", "on-blocks"])
@pytest.mark.parametrize("world,blocks,level", [(2, "2,1", 8), (2, "1,2", 7), (4, "2,2", 8), (8, "4,2", 8)])
def test_optical_flow_on_several_blocks(world, blocks, level, variant):
    """Benchmark/OptFlow2D on several blocks.  default: fused wavefront sweeps over the block
    plus the neighbours' layers, coarsest level agglomerated onto rank 0 for the coupled CG, the cycle one CUDA graph;
    on-blocks: colour kernels with `communicate` in between and the CG on the blocks with global sums.  Flow field
    bit-identical to the same program on one block and to the oracle, norms within 1e-12, CG iteration counts equal
    (tests/mgpu/optflow_worker.py)."""
    if _ngpus() < world:
        pytest.skip("needs %d GPUs" % world)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", str(29780 + world), os.path.join(ROOT, "tests", "mgpu", "optflow_worker.py"),
           "--level", str(level), "--blocks", blocks]
    env = dict(os.environ)
    if variant == "on-blocks":
        env.update(EXAMG_OF_NO_AGG="1", EXAMG_OF_NO_FUSED_BLOCKS="1")
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0 and "OPTFLOW MGPU OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]


def test_large_hybrid_golden():
    """Testing/Large/Hybrid: 2560 x 1024 x 2048 cells (5.4e9 unknowns) on 4 GPUs must print the reference's golden
    stdout (the oracle cannot hold this size; the golden vector comes from the reference's 4 x 10-way hybrid run)."""
    if _ngpus() < 4:
        pytest.skip("needs 4 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "4",
           "--master-addr", "127.0.0.1", "--master-port", "29741", os.path.join(ROOT, "tests", "mgpu", "large_worker.py")]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0 and "LARGE GOLDEN OK" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
