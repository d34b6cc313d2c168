"""World-size-2 (or more) check of the host-side multi-GPU plumbing on CPU (gloo backend):
the wiring all-gather of exastencils_b200.distributed.wire_peers with a stand-in for libexamg's
examg_comm_* entry points (no GPU, no CUDA IPC), the block decomposition helpers and the
max-over-ranks reduction.  Launched by tests/test_distributed_cpu.py through torch.distributed.run."""
import ctypes as C
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


class FakeLib:
    """Mimics the contract of examg_comm_blob_size / export / import (include/examg.h)."""
    BLOB = 96

    def __init__(self, rank, world):
        self.rank, self.world, self.imported = rank, world, None

    def examg_comm_blob_size(self, h):
        return self.BLOB

    def examg_comm_export(self, h, blob):
        payload = struct.pack("<II", 0x45584D47, self.rank) + bytes([self.rank + 1]) * (self.BLOB - 8)
        C.memmove(blob, payload, self.BLOB)
        return 0

    def examg_comm_import(self, h, joined, world):
        if world != self.world or len(joined) != world * self.BLOB:
            return -1
        for r in range(world):
            magic, rk = struct.unpack_from("<II", joined, r * self.BLOB)
            if magic != 0x45584D47 or rk != r or joined[r * self.BLOB + 8] != r + 1:
                return -1
        self.imported = joined
        return 0


class FakeCtx:
    def __init__(self, rank, world):
        self.L, self.h = FakeLib(rank, world), None


def main():
    import torch.distributed as dist
    from exastencils_b200 import distributed as D
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    ctx = FakeCtx(rank, world)
    D.wire_peers(ctx, dist)
    assert ctx.L.imported is not None and len(ctx.L.imported) == world * FakeLib.BLOB
    # all-gather keeps rank order and content
    got = D.all_gather_bytes(dist, bytes([rank]) * 5)
    assert got == [bytes([r]) * 5 for r in range(world)]
    # per-block maxima -> global maximum
    assert D.max_over_ranks(dist, float(rank) + 0.5) == world - 0.5
    # decomposition: the blocks of all ranks tile the global grid, shared nodes included
    nb = D.BLOCK_GRIDS[world] if world in D.BLOCK_GRIDS else (world, 1, 1)
    gshape = tuple(nb[d] * 8 + 1 for d in (2, 1, 0))
    cover = np.zeros(gshape, dtype=np.int64)
    for r in range(world):
        cover[D.block_slices(r, nb, gshape)] += 1
    assert cover.min() == 1
    sl = D.block_slices(rank, nb, gshape)
    assert tuple(s.stop - s.start for s in sl) == (9, 9, 9)
    for (d, side), other in D.neighbours(rank, nb).items():
        assert D.neighbours(other, nb)[(d, 1 - side)] == rank
    dist.barrier()
    dist.destroy_process_group()
    print("CPU WORKER %d OK" % rank)


if __name__ == "__main__":
    main()
