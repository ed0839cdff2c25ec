"""world_size-2 Gloo tests of the data-parallel plumbing (no GPU): bucketed all-reduce of a flat gradient
buffer equals the sum over ranks, whatever the bucket size; accumulation (no_sync) defers the exchange."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mingpt_distributed_b200.parallel import GradBuckets


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, cap_mb, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    spans = [(0, 640), (640, 4096), (4096, 8192), (8192, 12288), (12288, 20000)]
    g = torch.Generator().manual_seed(10 + rank)
    flat = torch.randn(20000, generator=g)
    mine = flat.clone()
    gb = GradBuckets(flat, spans, bucket_cap_mb=cap_mb)
    # accumulation micro-step: nothing is exchanged
    gb.enabled = False
    for grp in range(len(spans) - 1, -1, -1):
        gb.on_group_done(grp)
    assert torch.equal(flat, mine) and gb.launched == 0
    gb.enabled = True
    for grp in range(len(spans) - 1, -1, -1):     # the order backward finishes the layer groups
        gb.on_group_done(grp)
    gb.finish()
    others = [torch.randn(20000, generator=torch.Generator().manual_seed(10 + r)) for r in range(world)]
    want = sum(others)
    q.put((rank, torch.allclose(flat, want, atol=1e-6), gb.launched, len(gb.buckets), gb.grad_scale))
    dist.destroy_process_group()


def _run(cap_mb):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, cap_mb, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return res


def test_bucketed_allreduce_small_buckets():
    for rank, ok, launched, nb, scale in _run(cap_mb=0.01):
        assert ok and launched == nb and nb >= 3 and scale == 0.5


def test_bucketed_allreduce_single_bucket():
    for rank, ok, launched, nb, scale in _run(cap_mb=1000.0):
        assert ok and launched == 1 and nb == 1
