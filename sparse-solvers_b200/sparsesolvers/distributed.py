"""Multi-GPU layer of the batched solve (new functionality: the reference is single-process).

Signals are independent, so the batch shards with no data-path collective:
  - rank 0 owns the dictionary; it is replicated once with a broadcast (NCCL over NVLink on GPUs),
  - every rank solves its contiguous slice [lo, hi) of the batch against its replica,
  - the per-signal reports (and, when asked, the coefficients) are gathered to rank 0.
Nothing is exchanged between iterations.

The functions take the solve callable as an argument so the same code path runs with the `gloo`
backend on CPU tensors in the tests (tests/test_distributed_cpu.py) and with `nccl` in bench.py.
"""
import numpy as np


def shard_bounds(batch, world_size, rank):
    """Contiguous, balanced split of `batch` signals: the first (batch % world) ranks get one more."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world size")
    base, rem = divmod(int(batch), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def broadcast_dictionary(A, dist, device=None, src=0):
    """Replicates rank `src`'s dictionary on every rank.  A: torch tensor on `src` (any tensor of the right
    shape/dtype elsewhere); returns the replica (on `device` when given)."""
    import torch
    t = A if isinstance(A, torch.Tensor) else torch.as_tensor(np.ascontiguousarray(A))
    if device is not None:
        t = t.to(device)
    t = t.contiguous()
    dist.broadcast(t, src=src)
    return t


def gather_reports(iters, errs, dist, dst=0):
    """Gathers per-rank (iters, errs) tensors to `dst`, in rank order.  Shards may differ in length by one."""
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    n = torch.tensor([iters.numel()], dtype=torch.int64, device=iters.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    cap = max(sizes) if sizes else 0
    packed = torch.zeros((cap, 2), dtype=torch.float64, device=iters.device)
    packed[: iters.numel(), 0] = iters.to(torch.float64)
    packed[: iters.numel(), 1] = errs.to(torch.float64)
    out = [torch.zeros_like(packed) for _ in range(world)]
    dist.all_gather(out, packed)
    if rank != dst:
        return None, None
    it = torch.cat([o[:s, 0] for o, s in zip(out, sizes)]).to(torch.int64)
    er = torch.cat([o[:s, 1] for o, s in zip(out, sizes)])
    return it, er


def gather_coefficients(X, dist, dst=0):
    """Gathers the (rows, N) coefficient shards to `dst` in rank order (dense; 4*N bytes per signal)."""
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    n = torch.tensor([X.shape[0]], dtype=torch.int64, device=X.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    cap = max(sizes)
    pad = X if X.shape[0] == cap else torch.cat([X, X.new_zeros((cap - X.shape[0], X.shape[1]))])
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad.contiguous())
    if rank != dst:
        return None
    return torch.cat([o[:s] for o, s in zip(out, sizes)])


def solve_sharded(solve_fn, Y_full_or_shard, batch, dist, already_sharded=False, gather_x=False):
    """Runs `solve_fn(Y_shard) -> (X, iters, errs)` (torch tensors) on this rank's slice and gathers the reports.

    Y_full_or_shard: the full (batch, M) signal tensor (every rank slices its part) or, with
    already_sharded=True, this rank's shard.  Returns (X_shard, iters_all, errs_all, X_all) where the
    *_all entries are None except on rank 0 (X_all only when gather_x)."""
    world, rank = dist.get_world_size(), dist.get_rank()
    if already_sharded:
        Ys = Y_full_or_shard
    else:
        lo, hi = shard_bounds(batch, world, rank)
        Ys = Y_full_or_shard[lo:hi]
    X, it, er = solve_fn(Ys)
    it_all, er_all = gather_reports(it, er, dist)
    X_all = gather_coefficients(X, dist) if gather_x else None
    return X, it_all, er_all, X_all
