"""Host-side logic of the points-sharded data-parallel path, world_size 2 over gloo on CPU."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from deepearth_b200.distributed import allreduce_table_grads, flat_grad_buffers, shard_bounds


def test_shard_bounds_cover_batch():
    for n in (0, 1, 7, 2 ** 26, 2 ** 22 + 3):
        for w in (1, 2, 3, 4, 8):
            spans = [shard_bounds(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # two "tables" whose grads are views of one flat buffer, as Grid4DFunction produces
        p1 = torch.nn.Parameter(torch.zeros(8, 2))
        p2 = torch.nn.Parameter(torch.zeros(4, 2))
        flat = torch.arange(24, dtype=torch.float32) * (rank + 1)
        p1.grad = flat[:16].view(8, 2)
        p2.grad = flat[16:].view(4, 2)
        bufs = flat_grad_buffers([p1, p2])
        assert len(bufs) == 1 and bufs[0].numel() == 24
        allreduce_table_grads([p1, p2], average=True)
        want = torch.arange(24, dtype=torch.float32) * (sum(range(1, world + 1)) / world)
        ok = torch.allclose(torch.cat([p1.grad.flatten(), p2.grad.flatten()]), want)
        # level runs of the overlapped backward / all-reduce: contiguous, ordered, covering
        from deepearth_b200.distributed import GradSync
        runs = GradSync(chunks=4).level_runs(64)
        ok = ok and runs == [(0, 16), (16, 32), (32, 48), (48, 64)] and GradSync(chunks=8).level_runs(3) == [(0, 1), (1, 2), (2, 3)]
        b, e = shard_bounds(1000, rank, world)
        tot = torch.tensor([e - b])
        dist.all_reduce(tot)
        q.put((rank, bool(ok), int(tot)))
    finally:
        dist.destroy_process_group()


def test_allreduce_flat_grads_world2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok and tot == 1000 for _, ok, tot in res)
