"""Development aid: config-3 training step (fwd + L1 + bwd + Adam, batch 16 x 128x128) eager vs
captured in one CUDA graph.  Prints CPU enqueue time, GPU time per step for both, and checks that
graph replays follow the eager trajectory (same parameters after N steps)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import bench
from transmar_gan_b200 import dp, train

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
S = int(sys.argv[2]) if len(sys.argv) > 2 else 128
g = torch.Generator().manual_seed(999)
x = (torch.rand(B, 1, S, S, generator=g) * 2 - 1).cuda()
t = (torch.rand(B, 1, S, S, generator=g) * 2 - 1).cuda()


def make(graph):
    torch.manual_seed(0)
    m = bench.build_model().train()
    return m, train.L1TrainStep(m, lr=1e-4, betas=(0.5, 0.999), graph=graph)


def timeit(step, n):
    for _ in range(3):
        step(x, t)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    c0 = time.perf_counter(); e0.record()
    for _ in range(n):
        loss = step(x, t)
    e1.record(); c1 = time.perf_counter()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, (c1 - c0) * 1e3 / n, float(loss)


m_e, step_e = make(False)
gpu_ms, cpu_ms, loss = timeit(step_e, 10)
print(f"eager : {gpu_ms:.3f} ms/step on the device, {cpu_ms:.3f} ms/step CPU enqueue, loss {loss:.6f}")
m_g, step_g = make(True)
gpu_ms, cpu_ms, loss = timeit(step_g, 10)
print(f"graph : {gpu_ms:.3f} ms/step on the device, {cpu_ms:.3f} ms/step CPU enqueue, loss {loss:.6f}")
worst = 0.0
for (n, a), (_, b) in zip(m_e.named_parameters(), m_g.named_parameters()):
    d = (a - b).abs().max().item() / (a.abs().max().item() + 1e-12)
    worst = max(worst, d)
print(f"max relative parameter difference after 13 steps, eager vs graph: {worst:.3e}")
