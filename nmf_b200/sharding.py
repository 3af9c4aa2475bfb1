"""1-D sharding of V across the GPUs of one box (MPI-FAUN layout).

Rank p holds the row shard V[I_p, :] (for the W update) and the column shard V[:, J_p] (stored
transposed, for the H update); both factors are replicated.  Per outer iteration each rank updates
its window W[I_p, :] and H[:, J_p] and the windows are all-gathered -- the only data-path exchange
besides the scalar objective.  Everything here is plumbing over torch.distributed and works on CPU
tensors too (gloo), which is how the host logic is tested without GPUs.
"""
from __future__ import annotations

from typing import List, Tuple

import torch
import torch.distributed as dist


def round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


def shard_bounds(L: int, world: int, align: int = 256) -> Tuple[int, int, List[Tuple[int, int]]]:
    """Padded length (a multiple of align * world), rows per rank, and each rank's [lo, hi) window."""
    Lp = round_up(max(L, 1), align * world)
    per = Lp // world
    return Lp, per, [(p * per, (p + 1) * per) for p in range(world)]


def live_rows(L: int, lo: int, hi: int) -> int:
    """True (unpadded) rows inside the window [lo, hi)."""
    return max(0, min(hi, L) - lo)


def allgather_windows(master: torch.Tensor, per: int, group=None) -> None:
    """master: [KP][world * per]; on entry rank p's window master[:, p*per:(p+1)*per] is current,
    on exit the whole tensor is.  One all-gather of KP * per elements per rank."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    KP = master.shape[0]
    assert master.shape[1] == world * per
    send = master[:, rank * per:(rank + 1) * per].contiguous()
    recv = torch.empty((world, KP, per), dtype=master.dtype, device=master.device)
    try:
        dist.all_gather_into_tensor(recv, send, group=group)
    except (RuntimeError, NotImplementedError):          # backends without the tensor variant
        parts = [torch.empty_like(send) for _ in range(world)]
        dist.all_gather(parts, send, group=group)
        recv = torch.stack(parts)
    master.view(KP, world, per).copy_(recv.permute(1, 0, 2))


def allgather_flat(full: torch.Tensor, mine: torch.Tensor, group=None) -> None:
    """full: contiguous 1-D tensor of world * len(mine) elements; rank p contributes `mine` as its p-th
    segment (the norm partials of the AO-ADMM residual test)."""
    world = dist.get_world_size(group)
    assert full.is_contiguous() and full.numel() == world * mine.numel()
    try:
        dist.all_gather_into_tensor(full, mine, group=group)
    except (RuntimeError, NotImplementedError):
        parts = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine, group=group)
        full.copy_(torch.cat(parts))


def broadcast_from_rank0(arr, shape, device, group=None):
    """fp64 host array of `shape`, the one rank 0 passes (the others pass None), on every rank.  The transfer
    runs on `device` tensors (NCCL needs CUDA memory, gloo takes the CPU)."""
    import numpy as np
    rank = dist.get_rank(group)
    src = dist.get_global_rank(group, 0) if group is not None else 0
    if rank == 0:
        t = torch.from_numpy(np.ascontiguousarray(arr, dtype=np.float64)).to(device)
        assert tuple(t.shape) == tuple(shape)
    else:
        t = torch.empty(tuple(shape), dtype=torch.float64, device=device)
    dist.broadcast(t, src=src, group=group)
    return t.cpu().numpy()


def allreduce_sum_(t: torch.Tensor, group=None) -> None:
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)


def allreduce_max_(t: torch.Tensor, group=None) -> None:
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)


def allreduce_min_key_(key: torch.Tensor, group=None) -> None:
    """MIN over ranks of order-preserving uint32 keys held in an int32 tensor (the min(V) slot of the ingest
    statistics): flipping the top bit maps unsigned order onto the signed order all_reduce compares in."""
    flipped = key ^ torch.tensor(-2 ** 31, dtype=torch.int32, device=key.device)
    dist.all_reduce(flipped, op=dist.ReduceOp.MIN, group=group)
    key.copy_(flipped ^ torch.tensor(-2 ** 31, dtype=torch.int32, device=key.device))


def exchange_blocks(send: List[torch.Tensor], recv: List[torch.Tensor], group=None) -> None:
    """All-to-all of unequal blocks: send[p] goes to rank p, recv[p] arrives from rank p (contiguous tensors;
    the rank's own block is copied).  Point-to-point operations, so that it runs on NCCL (NVLink) and gloo."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    recv[rank].copy_(send[rank])
    ops = []
    for p in range(world):
        if p == rank:
            continue
        peer = dist.get_global_rank(group, p) if group is not None else p
        ops.append(dist.P2POp(dist.irecv, recv[p], peer, group))
        ops.append(dist.P2POp(dist.isend, send[p], peer, group))
    if ops:
        for work in dist.batch_isend_irecv(ops):
            work.wait()
