"""World-size-2 gloo tests of the multi-GPU layer's host logic (sparsesolvers/distributed.py) on CPU tensors.
The per-rank solve is stood in for by the CPU oracle (test infrastructure): what is under test is the split,
the dictionary broadcast and the gathers, not the solver."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from sparsesolvers import distributed as ssd


def test_shard_bounds_cover_the_batch():
    for batch in (0, 1, 7, 8, 65536, 65537):
        for world in (1, 2, 3, 8):
            spans = [ssd.shard_bounds(batch, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == batch
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        ssd.shard_bounds(4, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, batch, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import port as oracle
    from sparsesolvers import synth
    rng_A = synth.dictionary(32, 96, dtype=np.float64, seed=3)
    A0 = torch.from_numpy(rng_A) if rank == 0 else torch.zeros((32, 96), dtype=torch.float64)
    A = ssd.broadcast_dictionary(A0, dist)
    Y, _ = synth.sparse_signals(rng_A, batch, 3, seed=4)     # every rank can regenerate the full batch
    Yt = torch.from_numpy(Y)

    def solve(Ys):
        X, it, er = oracle.homotopy(A.numpy(), Ys.numpy(), 1e-3, 20) if len(Ys) else (np.zeros((0, 96)), np.zeros(0, np.uint32), np.zeros(0))
        return torch.from_numpy(X), torch.from_numpy(it.astype(np.int64)), torch.from_numpy(er)

    X, it_all, er_all, X_all = ssd.solve_sharded(solve, Yt, batch, dist, gather_x=True)
    if rank == 0:
        Xr, itr, err = oracle.homotopy(rng_A, Y, 1e-3, 20)
        out.put((bool(torch.equal(A, torch.from_numpy(rng_A))), it_all.numpy().tolist() == itr.tolist(),
                 bool(np.allclose(er_all.numpy(), err)), bool(np.array_equal(X_all.numpy(), Xr))))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("batch", [9, 2])
def test_sharded_solve_matches_single_process(batch):
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, batch, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = out.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == (True, True, True, True)
