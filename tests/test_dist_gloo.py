"""world_size-2 checks of the row-sharded (one process per GPU) path on CPU with the gloo backend:
shard boundaries, the packed [H | g | obj] all-reduce that GLM.fused_stats ends with, and that every
rank reaches the same coefficients as the single-process oracle without any broadcast."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import load_golden
from nums_b200 import hostlogic as hl
from oracle import glm as oglm


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


class _CpuBlock:
    """Stands in for a DeviceBlock: GLM._allreduce only touches `.t`."""

    def __init__(self, t):
        self.t = t


def _worker(rank, world, port, X, y, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from nums_b200.glms import GLM
        model = object.__new__(GLM)          # no device needed for the collective plumbing
        model._sharded = None                # default: all-reduce iff a process group with world > 1 exists
        n, d = X.shape
        p = d + 1
        lo, hi = hl.row_shard(n, world, rank)
        local = oglm.BlockedGLM("logistic", [X[lo:hi]], [y[lo:hi]], "l2", 0.1)
        beta = np.zeros(p)
        iters = 0
        for _ in range(100):
            fwd = local.forward(beta)
            # local statistics WITHOUT penalties, packed like nb_glm_stats does
            _, mus = fwd
            w = mus[0] * (1.0 - mus[0])
            Xa = np.hstack([X[lo:hi], np.ones((hi - lo, 1))])
            H = Xa.T @ (w[:, None] * Xa)
            g = Xa.T @ (mus[0] - y[lo:hi])
            obj = np.sum(local.fam.objective_terms(y[lo:hi], mus[0], fwd[0][0]))
            packed = torch.from_numpy(np.concatenate([H.reshape(-1), g, [obj]]))
            total = model._allreduce(_CpuBlock(packed)).t.numpy()
            H, g = total[: p * p].reshape(p, p), total[p * p: p * p + p]
            g = g + local.l2vec * beta                      # penalties once, after the reduction
            H = H + np.diag(local.l2vec)
            beta = beta - np.linalg.inv(H) @ g
            iters += 1
            if np.max(np.abs(g)) <= 1e-4:
                break
        out[rank] = (beta, iters)
    finally:
        dist.destroy_process_group()


def test_sharded_newton_two_ranks_gloo():
    g = load_golden("glm_readme")
    X, y = g["X"][:6001], g["y"][:6001].astype(np.float64)     # odd row count: ragged last shard
    want, it = oglm.fit("logistic", X, y, 3001, solver="newton", penalty="l2", lam=0.1)
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), X, y, out), nprocs=2, join=True)
    assert set(out.keys()) == {0, 1}
    b0, i0 = out[0]
    b1, i1 = out[1]
    np.testing.assert_array_equal(b0, b1)          # identical on every rank: no broadcast needed
    assert i0 == i1 == it
    assert np.max(np.abs(b0 - want)) < 1e-8


def test_row_shards_cover_baseline_config():
    spans = [hl.row_shard(10 ** 8, 8, r) for r in range(8)]
    assert spans[0] == (0, 12_500_000) and spans[-1] == (87_500_000, 10 ** 8)
