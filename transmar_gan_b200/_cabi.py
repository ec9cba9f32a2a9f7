"""ctypes binding of libb200mar.so (the C ABI declared in include/b200mar.h).

There is no CPU implementation and no fallback: if the shared library is missing, or a call
returns non-zero, a RuntimeError is raised.  `build()` compiles the library in-tree with nvcc
for sm_100a (cross-compiles without a GPU).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
# B200MAR_LIB: development override (e.g. a second build with -DB200MAR_ATTN_PROF for the phase profile);
# the product library is always transmar_gan_b200/libb200mar.so
LIB_PATH = os.environ.get("B200MAR_LIB") or os.path.join(_HERE, "libb200mar.so")
INCLUDE = os.path.join(os.path.dirname(_HERE), "include")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]

_P = ctypes.c_void_p
_I = ctypes.c_int
_L = ctypes.c_longlong
_F = ctypes.c_float

# name -> argtypes ; every function returns int unless listed in _RESTYPES
SIGNATURES = {
    "b200mar_abi_version": [],
    "b200mar_last_error": [],
    "b200mar_device_ok": [_I],
    "b200mar_unigram_fwd": [_P] * 4 + [_I] * 3 + [_P],
    "b200mar_unigram_bwd": [_P] * 6 + [_I] * 3 + [_P],
    "b200mar_unigram_bwd_add": [_P] * 7 + [_I] * 3 + [_P],
    "b200mar_ngram_ctx_fwd": [_P] * 10 + [_I] * 4 + [_P],
    "b200mar_unigram_qkv_fwd": [_P] * 6 + [_I] * 4 + [_P],
    "b200mar_unigram_qkv_u_fwd": [_P] * 7 + [_I] * 4 + [_P],
    "b200mar_ngram_ctx_pre_fwd": [_P] * 8 + [_I] * 4 + [_P],
    "b200mar_ngram_ctx_infer_noalloc": [_P] * 12 + [_I] * 4 + [_P],
    "b200mar_ngram_ctx_infer_noalloc_max_windows": [],
    "b200mar_ngram_ctx_bwd": [_P] * 19 + [_I] * 4 + [_P],
    "b200mar_attn_fwd": [_P] * 11 + [_I] * 5 + [_P],
    "b200mar_attn_tc_fwd": [_P] * 11 + [_I] * 6 + [_P],
    "b200mar_attn_bwd": [_P] * 21 + [_I] * 5 + [_P],
    "b200mar_attn_bwd_tc": [_P] * 23 + [_I] * 5 + [_P],
    "b200mar_attn_tc_fwd_save": [_P] * 12 + [_I] * 5 + [_P],
    "b200mar_attn_tc_fwd_comp": [_P] * 12 + [_I] * 5 + [_P],
    "b200mar_mlp_fwd": [_P] * 10 + [_L, _P],
    "b200mar_mlp_bwd": [_P] * 15 + [_L, _P],
    "b200mar_mlp_bwd_tc": [_P] * 16 + [_L, _P],
    "b200mar_wgrad_tc": [_P] * 4 + [_L, _I, _I, _P],
    "b200mar_mlp_tc_fwd": [_P] * 11 + [_I, _L, _P],
    "b200mar_mlp_tma_fwd": [_P] * 9 + [_L, _P],
    "b200mar_mlp_tma_fwd_comp": [_P] * 9 + [_L, _P],
    "b200mar_final_recon_bwd": [_P] * 19 + [_I] * 5 + [_P],
    "b200mar_cascade_proj_bwd": [_P] * 11 + [_I] * 3 + [_P],
    "b200mar_bottleneck_bwd": [_P] * 23 + [_I] * 3 + [_P],
    "b200mar_shallow_tok_bwd": [_P] * 6 + [_I] * 3 + [_P],
    "b200mar_cast_stream": [_P, _P, _L, _I, _P],
    "b200mar_unigram_qkv_fwd_s16": [_P] * 6 + [_I] * 4 + [_P],
    "b200mar_attn_tc_fwd_s16": [_P] * 11 + [_I] * 5 + [_P],
    "b200mar_mlp_tma_fwd_s16": [_P] * 9 + [_L, _P],
    "b200mar_tc_error": [],
    "b200mar_patch_merge_fwd": [_P] * 5 + [_I] * 3 + [_P],
    "b200mar_patch_merge_tc_fwd": [_P] * 5 + [_I] * 3 + [_P],
    "b200mar_patch_merge_bwd": [_P] * 9 + [_I] * 3 + [_P],
    "b200mar_ln64_fwd": [_P] * 4 + [_L, _P],
    "b200mar_ln64_bwd": [_P] * 6 + [_L, _P],
    "b200mar_metal_losses": [_P] * 5 + [_I] * 4 + [_F] * 6 + [_P],
    "b200mar_radon_fwd": [_P] * 3 + [_I] * 5 + [_F, _P],
    "b200mar_radon_bwd": [_P] * 3 + [_I] * 5 + [_F, _P],
    "b200mar_pixel_shuffle_cat": [_P] * 4 + [_I] * 4 + [_P],
    "b200mar_shallow_tok_fwd": [_P] * 4 + [_I] * 3 + [_P],
    "b200mar_cascade_proj_fwd": [_P] * 7 + [_I] * 3 + [_P],
    "b200mar_cascade_proj_tc_fwd": [_P] * 7 + [_I] * 3 + [_P],
    "b200mar_bottleneck_cat_fwd": [_P] * 7 + [_I] * 3 + [_P],
    "b200mar_bottleneck_tail_fwd": [_P] * 9 + [_I] * 3 + [_P],
    "b200mar_bottleneck_tail_tc_fwd": [_P] * 9 + [_I] * 3 + [_P],
    "b200mar_final_recon_fwd": [_P] * 10 + [_I] * 5 + [_P],
    "b200mar_partition_index": [_P] * 2 + [_I] * 4 + [_P],
    "b200mar_shift_mask": [_P] + [_I] * 3 + [_P],
    "b200mar_rel_index": [_P, _P],
}
_RESTYPES = {"b200mar_last_error": ctypes.c_char_p}

_lib = None


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def needs_build() -> bool:
    if not os.path.isfile(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(INCLUDE, "b200mar.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    """nvcc -gencode arch=compute_100a,code=sm_100a … -shared → transmar_gan_b200/libb200mar.so"""
    if not force and not needs_build():
        return LIB_PATH
    tag = os.path.splitext(os.path.basename(LIB_PATH))[0].replace("libb200mar", "")
    objdir = os.path.join(_HERE, "build" + tag)
    os.makedirs(objdir, exist_ok=True)

    def compile_one(src):
        obj = os.path.join(objdir, src[:-3] + ".o")
        cmd = ["nvcc", *NVCC_FLAGS, *os.environ.get("B200MAR_NVCC_EXTRA", "").split(), "-I", INCLUDE, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, sources()))
    cmd = ["nvcc", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB_PATH, *objs]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    global _lib
    _lib = None
    return LIB_PATH


def lib() -> ctypes.CDLL:
    """Load the shared library (never builds implicitly on a GPU box: the .so ships in-tree)."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). transmar_gan_b200 has no CPU or PyTorch fallback.")
        L = ctypes.CDLL(LIB_PATH)
        for name, argtypes in SIGNATURES.items():
            fn = getattr(L, name)          # AttributeError if the export is missing
            fn.argtypes = argtypes
            fn.restype = _RESTYPES.get(name, ctypes.c_int)
        _lib = L
    return _lib


def check(code: int, what: str):
    if code != 0:
        err = lib().b200mar_last_error()
        raise RuntimeError(f"{what} failed with code {code} "
                           f"({err.decode() if err else 'invalid argument'})")
