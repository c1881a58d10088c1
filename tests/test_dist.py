"""World-size-2 tests of the N>1 host logic on CPU (gloo): pattern sharding + the single
all-reduce of [logL, grad].  The per-rank evaluator here is the CPU oracle (the checker),
standing in for the per-rank GPU engine, which `-m gpu` tests cover."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from b200phylo.parallel import allreduce_sum_, rank_world, shard_bounds
from b200phylo.workloads import Workload, make_workload
from helpers import grad_err, oracle_c, rel_err


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert rank_world() == (rank, world)
        w = make_workload("hky_g4_flu", n_tips=30, n_patterns=301)
        lo, hi = shard_bounds(w.n_patterns, world, rank)
        ws = Workload("s", w.tree, w.patterns.slice(lo, hi), w.subst_model, w.site_model)
        l, g = oracle_c(ws)
        buf = torch.empty(1 + g.size, dtype=torch.float64)
        buf[0] = l
        buf[1:] = torch.from_numpy(g)
        allreduce_sum_(buf)
        if rank == 0:
            np.save(out_path, buf.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_sharded_allreduce_world2(tmp_path):
    out = str(tmp_path / "r.npy")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = np.load(out)
    w = make_workload("hky_g4_flu", n_tips=30, n_patterns=301)
    l, g = oracle_c(w)
    assert rel_err(got[0], l) < 1e-12
    assert grad_err(got[1:], g) < 1e-10


def test_allreduce_is_noop_without_group():
    buf = torch.arange(5, dtype=torch.float64)
    assert torch.equal(allreduce_sum_(buf.clone()), buf)
    assert rank_world() == (0, 1)
