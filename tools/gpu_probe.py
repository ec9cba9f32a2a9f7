import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from transmar_gan_b200 import _cabi, ops
L = _cabi.lib()
torch.manual_seed(0)
def tf32(x):
    return (x.view(torch.int32) & ~0x1FFF).view(torch.float32)
for variant in (0, 1):
  for mode in (0, 2):
    for (N, K, col0) in ((64, 64, 0), (16, 128, 0), (128, 16, 0), (16, 128, 16), (192, 64, 4), (64, 96, 0)):
        A = torch.randn(128, K, device="cuda")
        B = torch.randn(N, K, device="cuda")
        D = torch.full((128, N), float("nan"), device="cuda")
        rc = L.b200mar_tc_probe(A.data_ptr(), B.data_ptr(), D.data_ptr(), N, K, mode + 16 * variant, col0, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        ref = (tf32(A).double() @ tf32(B).double().T).float()
        print(f"variant {variant} mode {mode} N={N:3d} K={K:3d} col0={col0} rc={rc} err={(D-ref).abs().max().item():.2e} tc_error={ops.tc_error()}")
