"""Host-side multi-GPU logic on CPU: world_size-2 `gloo` process group.

Covers view / row sharding, the flat gradient all-reduce (sum and average, parameters without a
gradient on one rank) and the replica checksum.  No CUDA, no C-ABI compute calls.
"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pings_b200 import parallel as PP


def test_shard_rows_and_views_cover_everything():
    for n in (0, 1, 7, 115200, 10001):
        for ws in (1, 2, 3, 8):
            spans = [PP.shard_rows(n, r, ws) for r in range(ws)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(ws - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    views = list(range(8))
    for ws in (1, 2, 4, 8):
        got = sorted(v for r in range(ws) for v in PP.shard_views(views, r, ws))
        assert got == views
        assert all(len(PP.shard_views(views, r, ws)) == 8 // ws for r in range(ws))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, ws: int, port: int, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        torch.manual_seed(0)  # replicated parameters: same init on every rank
        a = torch.nn.Parameter(torch.randn(5, 3))
        b = torch.nn.Parameter(torch.randn(7))
        c = torch.nn.Parameter(torch.randn(2, 2))   # receives a gradient on rank 0 only
        g = torch.Generator().manual_seed(100 + rank)   # rank-local "views"
        x = torch.randn(4, 5, generator=g)
        loss = (x @ a).sum() + (b * (rank + 1)).sum()
        if rank == 0:
            loss = loss + (c ** 2).sum()
        loss.backward()
        local = [p.grad.clone() if p.grad is not None else torch.zeros_like(p) for p in (a, b, c)]
        red = PP.FlatGradAllReduce([a, b, c], average=True)
        assert red.nbytes == (15 + 7 + 4) * 4
        red()
        # expected: mean over ranks of the local gradients
        gathered = [[torch.zeros_like(t) for _ in range(ws)] for t in local]
        for t, gl in zip(local, gathered):
            dist.all_gather(gl, t)
        ok = True
        for p, gl in zip((a, b, c), gathered):
            ok &= torch.allclose(p.grad, torch.stack(gl).mean(0), rtol=0, atol=1e-6)
        # one SGD step with the reduced gradient keeps the replicas bit-identical ...
        with torch.no_grad():
            for p in (a, b, c):
                p -= 0.1 * p.grad
        ok &= PP.replicas_in_sync([a, b, c])
        # ... and a rank-local perturbation is detected
        with torch.no_grad():
            if rank == 1:
                a[0, 0] += 1e-3
        ok &= not PP.replicas_in_sync([a, b, c])
        # attached mode: grads are views of the flat buffer, autograd accumulates into it, no pack / scatter copies
        torch.manual_seed(1)
        a2 = torch.nn.Parameter(torch.randn(5, 3))
        b2 = torch.nn.Parameter(torch.randn(7))
        c2 = torch.nn.Parameter(torch.randn(2, 2))
        red2 = PP.FlatGradAllReduce([a2, b2, c2], average=True)
        for _ in range(2):   # second step: attach() must clear the previous step's sums
            red2.attach()
            l2 = (x @ a2).sum() + (b2 * (rank + 1)).sum() + (x @ a2).pow(2).sum()
            if rank == 0:
                l2 = l2 + (c2 ** 2).sum()
            l2.backward()
            ok &= a2.grad.data_ptr() == red2.views[0].data_ptr()
            loc2 = [p.grad.clone() for p in (a2, b2, c2)]
            red2()
            for p, t in zip((a2, b2, c2), loc2):
                gl = [torch.zeros_like(t) for _ in range(ws)]
                dist.all_gather(gl, t)
                ok &= torch.allclose(p.grad, torch.stack(gl).mean(0), rtol=0, atol=1e-6)
                ok &= p.grad.data_ptr() >= red2.flat.data_ptr()
        # multimem=True off CUDA: the exchange stays the collective (the NVSwitch kernel needs symmetric CUDA memory)
        a3 = torch.nn.Parameter(torch.ones(4) * (rank + 1))
        red3 = PP.FlatGradAllReduce([a3], average=False, multimem=True)
        ok &= red3.backend == "collective"
        a3.grad = torch.ones(4) * (rank + 1)
        red3()
        ok &= torch.equal(a3.grad, torch.full((4,), 3.0))
        # query-row sharding: the union of the shards reproduces the full result
        full = torch.arange(10, dtype=torch.float32) ** 2
        lo, hi = PP.shard_rows(10, rank, ws)
        part = torch.zeros(10)
        part[lo:hi] = full[lo:hi]
        dist.all_reduce(part)
        ok &= torch.equal(part, full)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_flat_grad_allreduce_gloo_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=100) for _ in procs)
    for p in procs:
        p.join(30)
        assert p.exitcode == 0
    assert res == [(0, True), (1, True)]
