"""CPU tests (gloo, world_size 2) of the data-parallel host logic: flat gradient bucket,
parameter broadcast, batch sharding.  The bucket is device-agnostic, so a toy module stands in
for the generator here (the generator's forward needs a GPU); the GPU box runs the same code
over NCCL in bench.py."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from transmar_gan_b200 import dp


def _toy(seed):
    torch.manual_seed(seed)
    return torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.GELU(), torch.nn.LayerNorm(16),
                               torch.nn.Linear(16, 4))


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
        model = _toy(seed=100 + rank)                 # ranks start different on purpose
        dp.broadcast_parameters(model, src=0)
        bucket = dp.FlatGradBucket(model.parameters())
        opt = torch.optim.Adam(model.parameters(), lr=1e-2, betas=(0.5, 0.999))
        g = torch.Generator().manual_seed(7)
        X = torch.randn(8, 8, generator=g)
        Y = torch.randn(8, 4, generator=g)
        xs, ys = dp.shard_batch(X, rank, world), dp.shard_batch(Y, rank, world)
        for _ in range(3):
            bucket.zero()
            torch.nn.functional.l1_loss(model(xs), ys).backward()
            bucket.all_reduce()
            opt.step()
        q.put((rank, [p.detach().numpy().copy() for p in model.parameters()], bucket.flat.numpy().copy()))
    finally:
        dist.destroy_process_group()


def test_flat_bucket_views_and_zero():
    m = _toy(0)
    b = dp.FlatGradBucket(m.parameters())
    assert b.param_numel == sum(p.numel() for p in m.parameters())
    assert b.numel == b.flat.numel() == sum(-(-p.numel() // 4) * 4 for p in m.parameters())
    m(torch.randn(3, 8)).sum().backward()
    covered = torch.zeros(b.numel, dtype=torch.bool)
    for p, off in zip(m.parameters(), b.offsets):  # grads accumulate INTO the flat buffer, views 16-byte aligned
        assert off % 4 == 0 and p.grad.data_ptr() == b.flat.data_ptr() + 4 * off
        assert torch.equal(p.grad.flatten(), b.flat[off:off + p.numel()])
        covered[off:off + p.numel()] = True
    assert b.flat.abs().sum() > 0 and b.flat[~covered].abs().sum() == 0     # the padding stays zero
    packed = dp.FlatGradBucket(_toy(0).parameters(), align=1)
    assert packed.numel == packed.param_numel
    b.zero()
    assert b.flat.abs().sum() == 0 and all(p.grad is not None for p in m.parameters())
    assert b.all_reduce() is None                 # single process: no-op
    for p in m.parameters():
        p.grad = None
    with pytest.raises(RuntimeError):
        b.zero()


def test_shard_batch():
    x = torch.arange(12).reshape(6, 2)
    assert torch.equal(dp.shard_batch(x, 1, 3), x[2:4])
    with pytest.raises(ValueError):
        dp.shard_batch(x, 0, 4)


def test_dp_world2_matches_single_process():
    """2 ranks with half the batch each + averaged flat bucket == 1 process with the full batch
    (L1 mean loss: the mean of the two half-batch gradients is the full-batch gradient)."""
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # single-process reference
    model = _toy(seed=100)
    opt = torch.optim.Adam(model.parameters(), lr=1e-2, betas=(0.5, 0.999))
    g = torch.Generator().manual_seed(7)
    X = torch.randn(8, 8, generator=g)
    Y = torch.randn(8, 4, generator=g)
    for _ in range(3):
        opt.zero_grad()
        torch.nn.functional.l1_loss(model(X), Y).backward()
        opt.step()
    (r0, p0, f0), (r1, p1, f1) = res
    assert (f0 == f1).all()                                   # identical averaged buckets
    for a, b, c in zip(p0, p1, model.parameters()):
        assert (a == b).all()                                 # replicas stay in lock-step
        assert torch.allclose(torch.from_numpy(a), c.detach(), atol=1e-6)   # == full-batch run
