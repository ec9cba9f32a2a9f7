"""CPU tests (single process and gloo world_size 2) of the host logic of train.L1TrainStep /
train.GANTrainStep in eager mode: recipe order, flat-bucket all-reduces, requires_grad toggling of the
discriminator.  A toy convolutional generator stands in for NGswin (whose forward needs a GPU); the
graph-replay path is covered on the GPU box (tests/test_gpu_train.py, bench.py)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn.functional as F

from transmar_gan_b200 import dp, gan, train


def _nets(seed):
    torch.manual_seed(seed)
    netG = torch.nn.Sequential(torch.nn.Conv2d(1, 4, 3, padding=1), torch.nn.Tanh(), torch.nn.Conv2d(4, 1, 3, padding=1))
    netD = gan.MultiScaleDiscriminator(2, 4, 3, 2, True)
    return netG, netD


def _data(rank=0, world=1, B=4):
    g = torch.Generator().manual_seed(5)
    ct = torch.rand(B, 1, 32, 32, generator=g) * 2 - 1
    real = torch.rand(B, 1, 32, 32, generator=g) * 2 - 1
    return dp.shard_batch(ct, rank, world), dp.shard_batch(real, rank, world)


def _reference_gan_steps(netG, netD, ct, real, n):
    """The recipe of training/train_combined.py:399-500 written out plainly (lambda_phys = 0)."""
    oG = torch.optim.Adam(netG.parameters(), lr=1e-4, betas=(0.5, 0.999))
    oD = torch.optim.Adam(netD.parameters(), lr=2e-4, betas=(0.5, 0.999))
    out = []
    for _ in range(n):
        netD.zero_grad()
        with torch.no_grad():
            fake = netG(ct)
        rp, fp = torch.cat([ct, real], 1), torch.cat([ct, fake], 1)
        rl, _ = netD(rp, return_features=False)
        fl, _ = netD(fp, return_features=False)
        lD = gan.hinge_d_loss(rl, fl)
        lD.backward()
        oD.step()
        netG.zero_grad()
        fake = netG(ct)
        fl, ff = netD(torch.cat([ct, fake], 1), return_features=True)
        with torch.no_grad():
            _, rf = netD(rp, return_features=True)
        w = gan.compute_weight_map(ct, 1.0, 5, 3.0, 0.6)
        M = gan.extract_metal_mask(ct, 0.6)
        lG = (0.1 * gan.hinge_g_loss(fl) + 10.0 * gan.feature_matching_loss(rf, ff)
              + 1.0 * gan.compute_metal_aware_loss(fake, real, ct, 1.0, 5, 3.0, 0.6)
              + 0.2 * gan.compute_metal_aware_edge_loss(fake, real, w) + 0.5 * gan.metal_consistency_loss(fake, real, M))
        lG.backward()
        oG.step()
        out.append((lD.item(), lG.item()))
    return out


def test_gan_step_eager_matches_written_out_recipe():
    ct, real = _data()
    nG, nD = _nets(3)
    ref = _reference_gan_steps(nG, nD, ct, real, 3)
    nG2, nD2 = _nets(3)
    step = train.GANTrainStep(nG2, nD2, graph=False)
    got = [tuple(float(v) for v in step(ct, real)) for _ in range(3)]
    for (a, b), (c, d) in zip(ref, got):
        assert abs(a - c) < 1e-5 * max(1, abs(a)) and abs(b - d) < 1e-5 * max(1, abs(b)), (ref, got)
    for p, q in zip(list(nG.parameters()) + list(nD.parameters()), list(nG2.parameters()) + list(nD2.parameters())):
        assert torch.allclose(p, q, rtol=1e-4, atol=1e-6)
    assert all(p.requires_grad for p in nD2.parameters())        # restored after the G step


def test_l1_step_eager():
    nG, _ = _nets(1)
    ct, real = _data()
    step = train.L1TrainStep(nG, graph=False)
    l0 = float(step(ct, real))
    for _ in range(20):
        l = float(step(ct, real))
    assert l < l0


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
        nG, nD = _nets(10 + rank)                      # ranks start different on purpose
        dp.broadcast_parameters(nG)
        dp.broadcast_parameters(nD)
        step = train.GANTrainStep(nG, nD, graph=False)
        ct, real = _data(rank, world)
        for _ in range(2):
            step(ct, real)
        q.put((rank, [p.detach().numpy().copy() for p in list(nG.parameters()) + list(nD.parameters())]))
    finally:
        dist.destroy_process_group()


def test_gan_step_world2_ranks_stay_in_lockstep():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=180) for _ in range(world)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for a, b in zip(res[0][1], res[1][1]):            # identical weights on both ranks after 2 steps
        assert (abs(a - b) <= 1e-6 + 1e-5 * abs(a)).all()
