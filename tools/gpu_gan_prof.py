"""Development aid: a few config-4 GAN steps (batch 16 x 128x128) through train.GANTrainStep with eager launches
and the full loss (fused image terms + physics term on the Radon kernels), for ncu launch lists."""
import math, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from transmar_gan_b200 import gan, radon, train
torch.backends.cudnn.benchmark = True
B, S = 16, 128
netG = bench.build_model().train()
torch.manual_seed(4321)
netD = gan.MultiScaleDiscriminator(2, 64, 5, 3, True).cuda().train()
netD.apply(gan.weights_init)
step = train.GANTrainStep(netG, netD, graph=False, projector=radon.Radon(S, torch.linspace(0, math.pi, 180)))
g = torch.Generator().manual_seed(999)
ct = (torch.rand(B, 1, S, S, generator=g) - 0.5)                      # |ct| < 0.5: metal only where the disc is
real = torch.rand(B, 1, S, S, generator=g) * 2 - 1
yy, xx = torch.meshgrid(torch.arange(S), torch.arange(S), indexing="ij")
ct[:, :, ((yy - 40) ** 2 + (xx - 70) ** 2) <= 36] = 1.0
ct, real = ct.cuda(), real.cuda()
for _ in range(4):
    step(ct, real)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.profiler.start()            # ncu --profile-from-start off: only the timed steps are listed
e0.record()
for _ in range(2):
    lD, lG = step(ct, real)
e1.record(); torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ms/step", e0.elapsed_time(e1) / 2, "loss_D", lD.item(), "loss_G", lG.item())
