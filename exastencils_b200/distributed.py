"""Multi-GPU plumbing: one process per GPU, `torch.distributed` carries the wiring blobs.

The data path never goes through torch.distributed: after `wire_peers` every `communicate`,
reduction and the gather/scatter of the agglomerated coarse levels run inside libexamg over
NVLink peer memory (csrc/examg_comm.cu).  What happens here is the analogue of the
`MPI_Allgather` in INTEGRATION.md section 5.

Block grid and rank mapping follow the reference: `domain_rect_numBlocks_{x,y,z}`, rank -> block
index x fastest (C/s/e/domain/ir/IR_InitGeneratedDomain.scala:41-47).
"""
import ctypes as C

import numpy as np

from . import runtime as rt

# domain_rect_numBlocks for 1, 2, 4, 8 ranks
BLOCK_GRIDS = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}


def block_index(rank, num_blocks):
    """(bx, by, bz) of a rank: x fastest."""
    bx = rank % num_blocks[0]
    by = (rank // num_blocks[0]) % num_blocks[1]
    bz = rank // (num_blocks[0] * num_blocks[1])
    return bx, by, bz


def block_rank(idx, num_blocks):
    return idx[0] + num_blocks[0] * (idx[1] + num_blocks[1] * idx[2])


def neighbours(rank, num_blocks):
    """{(dim, side): rank} of the face neighbours (side 0 = -, 1 = +), as
    C/s/e/communication/DefaultNeighbors.scala:50-60 numbers them (2*dim + side)."""
    idx = list(block_index(rank, num_blocks))
    out = {}
    for d in range(3):
        for side, step in ((0, -1), (1, 1)):
            j = list(idx)
            j[d] += step
            if 0 <= j[d] < num_blocks[d]:
                out[(d, side)] = block_rank(j, num_blocks)
    return out


def block_slices(rank, num_blocks, global_shape):
    """Slices (z, y, x) of a global node array that one block holds (shared nodes included)."""
    idx = block_index(rank, num_blocks)
    sl = []
    for d in (2, 1, 0):
        n_glob = global_shape[2 - d]
        cells = (n_glob - 1) // num_blocks[d]
        sl.append(slice(idx[d] * cells, (idx[d] + 1) * cells + 1))
    return tuple(sl)


def all_gather_bytes(dist, payload, device=None):
    """All-gather equal-sized byte strings through the process group (gloo: CPU tensors,
    nccl: tensors on `device`).  Returns the list of payloads in rank order."""
    import torch
    world = dist.get_world_size()
    backend = dist.get_backend()
    dev = torch.device("cpu") if backend == "gloo" else torch.device(device if device is not None else "cuda")
    mine = torch.frombuffer(bytearray(payload), dtype=torch.uint8).to(dev)
    sizes = torch.tensor([mine.numel()], dtype=torch.int64, device=dev)
    all_sizes = [torch.zeros_like(sizes) for _ in range(world)]
    dist.all_gather(all_sizes, sizes)
    if len({int(s.item()) for s in all_sizes}) != 1:
        raise rt.ExaMGError("wiring blobs differ in size across ranks")
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine)
    return [bytes(t.cpu().numpy().tobytes()) for t in out]


def wire_peers(ctx, dist, device=None):
    """examg_comm_export on every rank, all-gather, examg_comm_import.  Call after every field
    that takes part in `communicate` exists (i.e. after the solver object was created)."""
    n = rt.check(ctx.L.examg_comm_blob_size(ctx.h))
    blob = C.create_string_buffer(n)
    rt.check(ctx.L.examg_comm_export(ctx.h, blob))
    blobs = all_gather_bytes(dist, blob.raw, device)
    joined = b"".join(blobs)
    rt.check(ctx.L.examg_comm_import(ctx.h, joined, len(blobs)))
    dist.barrier()


def max_over_ranks(dist, value, device=None):
    """`examg_max_abs_diff` returns the per-block maximum; the global one is the max over blocks."""
    import torch
    backend = dist.get_backend()
    dev = torch.device("cpu") if backend == "gloo" else torch.device(device if device is not None else "cuda")
    t = torch.tensor([value], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
