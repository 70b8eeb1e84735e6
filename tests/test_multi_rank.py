"""CPU test of the N>1 path (world_size 2, gloo): contiguous sharding + ordered gather restore the single-rank result.
The per-shard compute is done by the oracle here (no GPU in this container); the sharding/gather code is the product's."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import casegen
from panvc3_b200.gather import gather_ordered, local_dict, shard_bounds


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, seed, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle_lib
        from parity import assert_same
        msa, b = casegen.random_batch(seed, 501, n_cols=400)
        lib = oracle_lib.load("port")
        idx = lib.make_index(*casegen.index_inputs(msa))
        lo, hi = shard_bounds(b.n, world, rank)
        local = idx.project_batch(b.slice(lo, hi), 0, -1)
        merged = gather_ordered(local_dict(local), device=torch.device("cpu"))
        if rank == 0:
            whole = idx.project_batch(b, 0, -1)
            assert_same(merged, whole, b, "gathered vs whole")
            assert int(whole.rr_len.sum()) > 0
            ret.put("ok")
        else:
            assert merged is None
    finally:
        dist.destroy_process_group()


def test_shard_bounds_cover_in_order():
    for n in (0, 1, 7, 1000):
        for w in (1, 2, 3, 8):
            b = [shard_bounds(n, w, r) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))


def test_two_rank_gather_restores_order():
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 4242, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert ret.get(timeout=5) == "ok"
