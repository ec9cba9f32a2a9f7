"""Development aid: exercise the tcgen05 operand conventions (tc_probe.cu) on a B200 and print the
error of each against a TF32-rounded matmul.  mode 0 = K-major B, 1 = MN-major B (4 candidate
LBO/SBO conventions), 2 = A operand from TMEM."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ctypes, subprocess
from transmar_gan_b200 import _cabi, ops
_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libtcprobe.so")
if not os.path.isfile(_SO) or os.path.getmtime(_SO) < os.path.getmtime(os.path.join(_HERE, "tc_probe.cu")):
    subprocess.check_call(["nvcc", *_cabi.NVCC_FLAGS, "-shared", "-I", _cabi.CSRC, "-I", _cabi.INCLUDE,
                           os.path.join(_HERE, "tc_probe.cu"), "-o", _SO])
L = ctypes.CDLL(_SO)
L.b200mar_tc_probe.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int] * 4 + [ctypes.c_void_p]
torch.manual_seed(0)


def tf32(x):
    return (x.view(torch.int32) & ~0x1FFF).view(torch.float32)


def run(N, K, mode, variant, col0):
    A = torch.randn(128, K, device="cuda")
    if mode == 1:
        B = torch.randn(K, N, device="cuda")          # MN-major source: [K][N]
        ref = (tf32(A).double() @ tf32(B).double()).float()
    else:
        B = torch.randn(N, K, device="cuda")
        ref = (tf32(A).double() @ tf32(B).double().T).float()
    D = torch.full((128, N), float("nan"), device="cuda")
    rc = L.b200mar_tc_probe(A.data_ptr(), B.data_ptr(), D.data_ptr(), N, K, mode + 16 * variant, col0,
                            torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    err = (D - ref).abs().max().item()
    print(f"mode {mode} variant {variant} N={N:3d} K={K:3d} col0={col0:3d} rc={rc} err={err:.2e} "
          f"tc_error={ops.tc_error()}", flush=True)


for (N, K, col0) in ((64, 64, 0), (16, 128, 0), (128, 16, 0), (16, 128, 16), (192, 64, 4), (48, 64, 48),
                     (16, 64, 224), (128, 16, 96)):
    run(N, K, 0, 0, col0)
for variant in (0, 1, 2, 3):
    for (N, K, col0) in ((16, 64, 0), (64, 64, 0), (16, 16, 0), (32, 8, 0)):
        run(N, K, 1, variant, col0)
for (N, K, col0) in ((64, 64, 0), (16, 64, 16)):
    run(N, K, 2, 0, col0)


def run16(N, K, mode, col0):
    A = torch.randn(128, K, device="cuda")
    B = torch.randn(N, K, device="cuda")
    ref = (A.half().double() @ B.half().double().T).float()
    D = torch.full((128, N), float("nan"), device="cuda")
    rc = L.b200mar_tc_probe(A.data_ptr(), B.data_ptr(), D.data_ptr(), N, K, mode, col0,
                            torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    print(f"f16 mode {mode} N={N:3d} K={K:3d} col0={col0:3d} rc={rc} err={(D - ref).abs().max().item():.2e} "
          f"tc_error={ops.tc_error()}", flush=True)


for mode in (4, 5):
    for (N, K, col0) in ((64, 64, 0), (48, 64, 48), (64, 32, 96), (32, 128, 16), (32, 64, 0), (128, 128, 8)):
        run16(N, K, mode, col0)


def run_bf16(N, K, mode, col0):
    """mode 6: D[m][n] = sum_k A[k][m] B[k][n] (both MN-major); mode 7: A [128][K] from TMEM, B [K][N] MN-major."""
    A = torch.randn(K, 128, device="cuda") if mode == 6 else torch.randn(128, K, device="cuda")
    B = torch.randn(K, N, device="cuda")
    Ab, Bb = A.bfloat16().double(), B.bfloat16().double()
    ref = ((Ab.T if mode == 6 else Ab) @ Bb).float()
    D = torch.full((128, N), float("nan"), device="cuda")
    rc = L.b200mar_tc_probe(A.data_ptr(), B.data_ptr(), D.data_ptr(), N, K, mode, col0,
                            torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    print(f"bf16 mode {mode} N={N:3d} K={K:3d} col0={col0:3d} rc={rc} err={(D - ref).abs().max().item():.2e} "
          f"tc_error={ops.tc_error()}", flush=True)


for mode in (6, 7):
    for (N, K, col0) in ((64, 128, 0), (64, 16, 0), (192, 128, 16), (128, 64, 32), (16, 32, 0), (256, 128, 0)):
        run_bf16(N, K, mode, col0)
