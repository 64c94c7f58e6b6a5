"""ctypes binding of liblora_b200.so (C ABI in include/lora_b200.h).

PyTorch is used for device memory and streams only: every call passes raw
device pointers and the current CUDA stream.  There is no CPU path -- if the
shared library is missing or fails to load, importing the product fails loudly.
"""
from __future__ import annotations

import ctypes
import functools
import os
from ctypes import c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_uint32, c_uint64, c_void_p

import torch

from ._build import LIB_PATH

DTYPE_BF16 = 0
DTYPE_FP16 = 1

_lib = None

_SIGNATURES = {
    "lora_last_error": (c_char_p, []),
    "lora_linear_rpad": (c_int32, [c_int32]),
    "lora_linear_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int64, c_int32]),
    "lora_linear_fwd": (
        c_int,
        [c_void_p] * 9 + [c_int64, c_int64, c_int64, c_int32, c_float, c_float, c_uint64, c_uint64, c_int32, c_int32,
                          c_void_p, c_size_t, c_void_p],
    ),
    "lora_linear_fwd_devrng": (
        c_int,
        [c_void_p] * 9 + [c_int64, c_int64, c_int64, c_int32, c_float, c_float, c_uint64, c_uint64, c_void_p, c_int32, c_int32,
                          c_void_p, c_size_t, c_void_p],
    ),
    "lora_linear_bwd": (
        c_int,
        [c_void_p] * 10 + [c_int64, c_int64, c_int64, c_int32, c_float, c_float, c_int32, c_int32, c_int32, c_int32,
                           c_void_p, c_size_t, c_void_p],
    ),
    "lora_linear_base": (c_int, [c_void_p] * 4 + [c_int64, c_int64, c_int64, c_int32, c_int32, c_void_p]),
    "lora_linear_grad_b_static": (c_int, [c_void_p] * 3 + [c_int64, c_int64, c_int32, c_float, c_int32, c_int32,
                                                           c_void_p]),
    "lora_frozen_linear_fwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64,
                                       c_int32, c_int32, c_void_p]),
    "lora_frozen_linear_bwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int32,
                                       c_int32, c_void_p]),
    "lora_frozen_linear_act_fwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_int64,
                                           c_int64, c_int64, c_int32, c_int32, c_int32, c_void_p]),
    "lora_act_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int32, c_int32, c_void_p]),
    "lora_p2p_buffer_bytes": (c_size_t, [c_int64]),
    "lora_p2p_alloc": (c_int, [c_int64, ctypes.POINTER(c_void_p), c_void_p]),
    "lora_p2p_open": (c_int, [c_void_p, ctypes.POINTER(c_void_p)]),
    "lora_p2p_close": (c_int, [c_void_p]),
    "lora_p2p_free": (c_int, [c_void_p]),
    "lora_bucket_allreduce_avg": (c_int, [c_void_p, c_int64, ctypes.POINTER(c_void_p), c_int32, c_int32, c_int64, c_int32,
                                          c_void_p]),
    "lora_merge_weight": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int32, c_float, c_int32, c_int32, c_void_p]),
    "lora_ce_fwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int32, c_void_p]),
    "lora_ce_bwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64,
                            c_int32, c_void_p]),
    "lora_debug_set_phase_mask": (None, [c_int32]),
    "lora_dropout_keep_mask": (c_int, [c_void_p, c_int64, c_int64, c_float, c_uint64, c_uint64, c_void_p]),
}


def exported_symbols():
    """Names declared in include/lora_b200.h (used by the CPU-side ABI test)."""
    return sorted(_SIGNATURES)


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "There is no CPU or PyTorch fallback for the fused LoRA kernels."
            )
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError here = ABI drift, fail loudly
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def last_error() -> str:
    return lib().lora_last_error().decode()


def _check(rc: int, what: str):
    if rc != 0:
        raise RuntimeError(f"{what}: {lib().lora_last_error().decode()}")


def dtype_code(dt: torch.dtype) -> int:
    if dt == torch.bfloat16:
        return DTYPE_BF16
    if dt == torch.float16:
        return DTYPE_FP16
    raise RuntimeError(f"fused LoRA kernels compute in bf16 or fp16 with fp32 accumulation; got {dt}")


def _ptr(t):
    return None if t is None else t.data_ptr()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream() -> int:
    """cudaStream_t of torch's current stream on the current device (the raw getter skips building a Stream object:
    this sits on the per-layer launch path, which bounds the small models' step time)."""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts):
    for t in ts:
        if t is not None and (not t.is_cuda or not t.is_contiguous()):
            raise RuntimeError("fused LoRA kernels need contiguous CUDA tensors (no CPU fallback)")


@functools.lru_cache(maxsize=None)
def rpad(r: int) -> int:
    return int(lib().lora_linear_rpad(int(r)))


@functools.lru_cache(maxsize=4096)
def workspace_bytes(M: int, K: int, N: int, r: int) -> int:
    return int(lib().lora_linear_workspace_bytes(M, K, N, r))


def linear_fwd(x2d, W, bias, A, B, scaling, p, seed, offset, w_is_KN, workspace, out=None, save_xd=False,
               offset_dev=None):
    """Y, H, keep[, Xd] = fused forward.  x2d [M,K] 16-bit, A [r,K] / B [N,r] fp32.
    save_xd: also return keep .* X (None when p == 0) for linear_bwd(..., x_is_dropped=True).
    offset_dev: optional int64/uint64 CUDA scalar added to `offset` on the device (CUDA-graph replay)."""
    _need_cuda(x2d, W, bias, A, B, workspace)
    M, K = x2d.shape
    N = W.shape[1] if w_is_KN else W.shape[0]
    r = A.shape[0]
    Y = out if out is not None else torch.empty((M, N), dtype=x2d.dtype, device=x2d.device)
    H = torch.empty((M, rpad(r)), dtype=x2d.dtype, device=x2d.device)
    keep = torch.empty((M, K // 8), dtype=torch.uint8, device=x2d.device) if p > 0 else None
    xd = torch.empty_like(x2d) if (save_xd and p > 0) else None
    if offset_dev is not None and (not offset_dev.is_cuda or offset_dev.element_size() != 8):
        raise RuntimeError("offset_dev must be a 64-bit integer CUDA tensor")
    rc = lib().lora_linear_fwd_devrng(
        _ptr(x2d), _ptr(W), _ptr(bias), _ptr(A), _ptr(B), _ptr(Y), _ptr(H), _ptr(keep), _ptr(xd), M, K, N, r,
        float(scaling),
        float(p), int(seed), int(offset), _ptr(offset_dev), dtype_code(x2d.dtype), int(bool(w_is_KN)), _ptr(workspace),
        workspace.numel() * workspace.element_size(), _stream(),
    )
    _check(rc, "lora_linear_fwd")
    if save_xd:
        return Y, H, keep, xd
    return Y, H, keep


def linear_bwd(dy2d, x2d, W, A, B, H, keep, scaling, p, w_is_KN, workspace, need_dx=True, dA=None, dB=None,
               accumulate=False, x_is_dropped=False):
    """dX (or None), dA [r,K] fp32, dB [N,r] fp32.  x2d is X, or forward's Xd with x_is_dropped."""
    _need_cuda(dy2d, x2d, W, A, B, H, keep, workspace)
    M, K = x2d.shape
    N = dy2d.shape[1]
    r = A.shape[0]
    dX = torch.empty_like(x2d) if need_dx else None
    if dA is None:
        dA = torch.empty((r, K), dtype=torch.float32, device=x2d.device)
    if dB is None:
        dB = torch.empty((N, r), dtype=torch.float32, device=x2d.device)
    rc = lib().lora_linear_bwd(
        _ptr(dy2d), _ptr(x2d), _ptr(W), _ptr(A), _ptr(B), _ptr(H), _ptr(keep), _ptr(dX), _ptr(dA), _ptr(dB), M, K, N,
        r, float(scaling), float(p), dtype_code(x2d.dtype), int(bool(w_is_KN)), int(bool(accumulate)),
        int(bool(x_is_dropped)), _ptr(workspace), workspace.numel() * workspace.element_size(), _stream(),
    )
    _check(rc, "lora_linear_bwd")
    return dX, dA, dB


def linear_base(x2d, W, bias, w_is_KN):
    """Y = X W^T + bias through the same tcgen05 kernel (adapter disabled / merged)."""
    _need_cuda(x2d, W, bias)
    M, K = x2d.shape
    N = W.shape[1] if w_is_KN else W.shape[0]
    Y = torch.empty((M, N), dtype=x2d.dtype, device=x2d.device)
    rc = lib().lora_linear_base(_ptr(x2d), _ptr(W), _ptr(bias), _ptr(Y), M, K, N, dtype_code(x2d.dtype),
                                int(bool(w_is_KN)), _stream())
    _check(rc, "lora_linear_base")
    return Y


PHASE_PACK, PHASE_ROWPROJ, PHASE_GEMM, PHASE_COLPROJ, PHASE_ALL = 1, 2, 4, 8, 15


def set_phase_mask(mask: int = PHASE_ALL):
    """Profiling only (see include/lora_b200.h): choose which kernels fwd/bwd launch."""
    lib().lora_debug_set_phase_mask(int(mask))


def dropout_keep_mask(M, K, p, seed, offset, device="cuda"):
    keep = torch.empty((M, K // 8), dtype=torch.uint8, device=device)
    rc = lib().lora_dropout_keep_mask(_ptr(keep), M, K, float(p), int(seed), int(offset), _stream())
    _check(rc, "lora_dropout_keep_mask")
    return keep
