"""CPU test of the N > 1 host logic (gloo, world_size 2): every rank evaluates its row shard, the
additive result block [lp, dalpha, dbeta...] is all-reduced, and the sum equals the unsharded
evaluation -- the property the NCCL path relies on.  The per-shard evaluation here is the plain-C
oracle (a checker standing in for the kernel; no GPU in this container)."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from math_b200.shard import rank_rows


def test_rank_rows_partition():
    for n in (0, 1, 7, 1000, 10_000_000):
        for world in (1, 2, 3, 8):
            blocks = [rank_rows(n, world, r) for r in range(world)]
            assert sum(b[1] for b in blocks) == n
            pos = 0
            for r0, rows in blocks:
                if rows:
                    assert r0 == pos
                pos += rows


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    from oracle import loader as L
    from tests.helpers import synth_glm
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s = synth_glm("poisson_log", 5001, 24, seed=3)
    r0, rows = rank_rows(5001, world, rank)
    x = np.asfortranarray(s["x"][r0:r0 + rows])
    part = L.oracle_glm("poisson_log", x, s["y"][r0:r0 + rows], s["alpha"], s["beta"])
    block = torch.from_numpy(np.concatenate([[part["lp"]], part["dalpha"], part["dbeta"]]))
    dist.all_reduce(block)
    if rank == 0:
        full = L.oracle_glm("poisson_log", s["x"], s["y"], s["alpha"], s["beta"])
        q.put((block.numpy(), np.concatenate([[full["lp"]], full["dalpha"], full["dbeta"]])))
    dist.destroy_process_group()


def test_row_shards_allreduce_to_the_full_evaluation():
    sock = socket.socket()
    sock.bind(("127.0.0.1", 0))
    port = sock.getsockname()[1]
    sock.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    got, ref = q.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert np.max(np.abs(got - ref) / np.maximum(np.abs(ref), 1e-6 * np.max(np.abs(ref)))) < 1e-11
