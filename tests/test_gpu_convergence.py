"""Does the tensor-core training recipe (fp16 / tf32 forward, bf16-operand backward: gradients 1-4 % off the
fp32 ones per tensor) TRAIN like the reference's fp32 arithmetic?  300 steps of the config-3
recipe (fwd + L1 + bwd + Adam betas (.5,.999), lr 1e-4: scripts/run_ablation_studies.py:1525-1535) from identical
weights (W1) on an identical stream of structured synthetic slices (oracle/phantoms.py): THREE runs on the
tensor-core kernels (CUDA-graph replay, the benchmarked configuration) and THREE on the fp32 FMA-pipe kernels.
Training is chaotic and every backward accumulates weight gradients with atomics, so two runs of the SAME arm
already differ (measured on B200: fp32-vs-fp32 25-step window means up to 43 % apart when one run takes a loss
bump); single runs cannot be compared.  What licenses quoting steps/s on the fast recipe is that the two arms are
indistinguishable as DISTRIBUTIONS, compared through spike-robust statistics (medians):
  * every run learns: median of its last 100 losses < 60 % of the first loss;
  * tail (median of the last 100 losses), median over the three runs of each arm: within 8 %;
  * 25-step window medians, median over runs: mean gap over the windows < 8 %;
  * held-out loss, median over runs: within 15 %;
  * the spread between the arms' medians is reported beside the spread inside the fp32 arm.
"""
import pytest
import torch
import torch.nn.functional as F

from oracle import ngswin_oracle as O
from oracle.phantoms import phantom_batch

pytestmark = pytest.mark.gpu
STEPS, WIN, NRUN, LR = 300, 25, 3, 1e-4


def _run(T, precision, graph, data, lr):
    from transmar_gan_b200 import train
    old = T.ops.set_precision(precision)
    old_tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = precision != "fp32"       # the stock glue convolutions follow the arm
    try:
        m = T.NGswin()
        m.load_state_dict(O.synthetic_state_dict("W1", seed=1), strict=True)
        m = m.cuda().train()
        step = train.L1TrainStep(m, lr=lr, betas=(0.5, 0.999), graph=graph)
        losses = [step(x, t) for x, t in data]
        losses = torch.stack(losses).cpu()
        m.eval()
        g = torch.Generator().manual_seed(4321)
        xv, tv = phantom_batch(8, 64, g)
        with torch.no_grad():
            held = F.l1_loss(m(xv.cuda()), tv.cuda()).item()
        assert T.ops.tc_error() == 0
        return losses, held
    finally:
        torch.backends.cudnn.allow_tf32 = old_tf32
        T.ops.set_precision(old)


def test_loss_trajectory_tensor_core_vs_fp32():
    import transmar_gan_b200 as T
    g = torch.Generator().manual_seed(2024)
    pool = [phantom_batch(4, 64, g) for _ in range(40)]
    data = [(pool[i % 40][0].cuda(), pool[i % 40][1].cuda()) for i in range(STEPS)]
    runs = {"tc": [_run(T, "tf32", True, data, LR) for _ in range(NRUN)],
            "fp": [_run(T, "fp32", False, data, LR) for _ in range(NRUN)]}
    first = runs["fp"][0][0][0].item()
    tails = {k: torch.tensor([l[-100:].median().item() for l, _ in v]) for k, v in runs.items()}
    held = {k: torch.tensor([h for _, h in v]) for k, v in runs.items()}
    wins = {k: torch.stack([l.view(-1, WIN).median(1).values for l, _ in v]).median(0).values for k, v in runs.items()}
    gaps = (wins["tc"] - wins["fp"]).abs() / wins["fp"]
    tail_gap = abs(tails["tc"].median() - tails["fp"].median()).item() / tails["fp"].median().item()
    held_gap = abs(held["tc"].median() - held["fp"].median()).item() / held["fp"].median().item()
    fp_spread = ((tails["fp"].max() - tails["fp"].min()) / tails["fp"].median()).item()
    print("[convergence] window medians fp32:", " ".join(f"{v:.4f}" for v in wins["fp"].tolist()))
    print("[convergence] window medians tc  :", " ".join(f"{v:.4f}" for v in wins["tc"].tolist()))
    print(f"[convergence] first loss {first:.4f}; tails fp32 {[round(v, 4) for v in tails['fp'].tolist()]} tensor-core "
          f"{[round(v, 4) for v in tails['tc'].tolist()]}; tail gap of medians {tail_gap:.3f} (spread inside fp32 "
          f"{fp_spread:.3f}); window gap mean {gaps.mean():.3f} worst {gaps.max():.3f}; held-out fp32 "
          f"{[round(v, 4) for v in held['fp'].tolist()]} tensor-core {[round(v, 4) for v in held['tc'].tolist()]}")
    for l, _ in runs["tc"] + runs["fp"]:
        assert abs(l[0].item() - first) < 1e-3 * first                       # same start
        assert l[-100:].median().item() < 0.6 * first                        # every run learns
    assert tail_gap < 0.08, tail_gap
    assert gaps.mean().item() < 0.08, gaps
    assert held_gap < 0.15, held_gap
