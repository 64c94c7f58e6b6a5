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
    "lora_linear_packed_bytes": (c_size_t, [c_int64, c_int64, c_int32]),
    "lora_linear_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int64, c_int32]),
    "lora_linear_pack": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_int64, c_int64, c_int32, c_float, c_int32,
                                 c_void_p]),
    "lora_linear_fwd": (
        c_int,
        [c_void_p] * 8 + [c_int64, c_int64, c_int64, c_int32, c_float, c_float, c_uint64, c_uint64, c_int32, c_int32,
                          c_void_p, c_size_t, c_void_p],
    ),
    "lora_linear_fwd_devrng": (
        c_int,
        [c_void_p] * 8 + [c_int64, c_int64, c_int64, c_int32, c_float, c_float, c_uint64, c_uint64, c_void_p, c_int32, c_int32,
                          c_void_p, c_size_t, c_void_p],
    ),
    "lora_linear_bwd": (
        c_int,
        [c_void_p] * 9 + [c_int64, c_int64, c_int64, c_int32, c_float, c_float, c_int32, c_int32, c_int32,
                          c_void_p, c_size_t, c_void_p],
    ),
    "lora_linear_base": (c_int, [c_void_p] * 4 + [c_int64, c_int64, c_int64, c_int32, c_int32, c_void_p]),
    "lora_linear_grad_b_static": (c_int, [c_void_p] * 3 + [c_int64, c_int64, c_int32, c_float, c_int32, c_int32,
                                                           c_void_p]),
    "lora_frozen_linear_fwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64,
                                       c_int32, c_int32, c_void_p]),
    "lora_frozen_linear_add_fwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_int64,
                                           c_int64, c_int64, c_int32, c_int32, c_void_p]),
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
    "lora_adamw_clip_step": (c_int, [c_void_p, c_int32, c_int64, c_float, c_float, c_float, c_float, c_float, c_float,
                                     c_void_p, c_void_p]),
    "lora_debug_set_phase_mask": (None, [c_int32]),
    "lora_dropout_keep_mask": (c_int, [c_void_p, c_int64, c_int64, c_float, c_uint64, c_uint64, c_void_p]),
}


def exported_symbols():
    """Names declared in include/lora_b200.h (used by the CPU-side ABI test)."""
    return sorted(_SIGNATURES)


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        path = os.environ.get("LORA_B200_LIB") or LIB_PATH  # LORA_B200_LIB: an experiment build (_build.build_variant)
        if not os.path.exists(path):
            raise RuntimeError(
                f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "There is no CPU or PyTorch fallback for the fused LoRA kernels."
            )
        handle = ctypes.CDLL(path)
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


def _stream(device=None) -> int:
    """cudaStream_t of torch's current stream on `device` (default: the current device).  The raw getter skips
    building a Stream object: this sits on the per-layer launch path, which bounds the small models' step time."""
    idx = torch.cuda.current_device() if device is None or device.index is None else device.index
    if _raw_stream is not None:
        return _raw_stream(idx)
    return torch.cuda.current_stream(idx).cuda_stream


def _need_cuda(*ts):
    """Contiguous CUDA tensors, all on ONE device, and that device must be the current one: the library
    launches on the current device's stream, so a model living on cuda:1 while cuda:0 is current would otherwise
    run device-0 kernels on device-1 memory.  Callers wrap launches in `device_guard` (below), which makes the
    operands' device current for the duration of the call."""
    dev = None
    for t in ts:
        if t is None:
            continue
        if not t.is_cuda or not t.is_contiguous():
            raise RuntimeError("fused LoRA kernels need contiguous CUDA tensors (no CPU fallback)")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"fused LoRA kernels need all operands on one device; got {dev} and {t.device}")
    if dev is not None and dev.index != torch.cuda.current_device():
        raise RuntimeError(f"operands live on {dev} but the current CUDA device is {torch.cuda.current_device()}: "
                           "wrap the call in binding.device_guard(tensor)")


class device_guard:
    """`with device_guard(t):` makes t's device current for the kernels launched inside (a no-op, and nearly
    free, when it already is).  Every autograd node and module of this package launches under it."""

    __slots__ = ("idx", "prev")

    def __init__(self, t):
        if not t.is_cuda:
            raise RuntimeError("fused LoRA kernels need contiguous CUDA tensors (no CPU fallback)")
        self.idx = t.device.index if t.device.index is not None else torch.cuda.current_device()
        self.prev = -1

    def __enter__(self):
        cur = torch.cuda.current_device()
        if cur != self.idx:
            self.prev = cur
            torch.cuda.set_device(self.idx)
        return self

    def __exit__(self, *exc):
        if self.prev >= 0:
            torch.cuda.set_device(self.prev)
        return False


@functools.lru_cache(maxsize=None)
def rpad(r: int) -> int:
    return int(lib().lora_linear_rpad(int(r)))


@functools.lru_cache(maxsize=4096)
def workspace_bytes(M: int, K: int, N: int, r: int) -> int:
    """Scratch of the BACKWARD (dH, arrival counters, split partial sums); the forward needs none."""
    return int(lib().lora_linear_workspace_bytes(M, K, N, r))


@functools.lru_cache(maxsize=4096)
def packed_bytes(K: int, N: int, r: int) -> int:
    return int(lib().lora_linear_packed_bytes(K, N, r))


def pack_adapters(A, B, scaling, dtype, out=None):
    """A [r,K], B [N,r] fp32 -> the 16-bit operand block forward and backward read."""
    _need_cuda(A, B, out)
    r, K = A.shape
    N = B.shape[0]
    if out is None:
        out = torch.empty(packed_bytes(K, N, r), dtype=torch.uint8, device=A.device)
    rc = lib().lora_linear_pack(_ptr(A), _ptr(B), _ptr(out), out.numel(), K, N, r, float(scaling), dtype_code(dtype),
                                _stream())
    _check(rc, "lora_linear_pack")
    return out


def linear_fwd(x2d, W, bias, A, B, scaling, p, seed, offset, w_is_KN, packed=None, out=None, offset_dev=None,
               r=None):
    """Y, H, keep, packed = fused forward.  x2d [M,K] 16-bit, A [r,K] / B [N,r] fp32.
    packed: uint8 block of packed_bytes(K, N, r) that receives the 16-bit adapter operands (allocated when None)
    and must be handed to linear_bwd; with A = B = None it must already hold them (pack_adapters) and `r` is given.
    offset_dev: optional int64/uint64 CUDA scalar added to `offset` on the device (CUDA-graph replay)."""
    _need_cuda(x2d, W, bias, A, B, packed)
    M, K = x2d.shape
    N = W.shape[1] if w_is_KN else W.shape[0]
    r = A.shape[0] if A is not None else int(r)
    if packed is None:
        if A is None:
            raise RuntimeError("linear_fwd: without A and B the packed operand block must be given")
        packed = torch.empty(packed_bytes(K, N, r), dtype=torch.uint8, device=x2d.device)
    Y = out if out is not None else torch.empty((M, N), dtype=x2d.dtype, device=x2d.device)
    H = torch.empty((M, rpad(r)), dtype=x2d.dtype, device=x2d.device)
    keep = torch.empty((M, K // 8), dtype=torch.uint8, device=x2d.device) if p > 0 else None
    if offset_dev is not None and (not offset_dev.is_cuda or offset_dev.element_size() != 8):
        raise RuntimeError("offset_dev must be a 64-bit integer CUDA tensor")
    rc = lib().lora_linear_fwd_devrng(
        _ptr(x2d), _ptr(W), _ptr(bias), _ptr(A), _ptr(B), _ptr(Y), _ptr(H), _ptr(keep), M, K, N, r, float(scaling),
        float(p), int(seed), int(offset), _ptr(offset_dev), dtype_code(x2d.dtype), int(bool(w_is_KN)), _ptr(packed),
        packed.numel(), _stream(),
    )
    _check(rc, "lora_linear_fwd")
    return Y, H, keep, packed


def linear_bwd(dy2d, x2d, W, packed, H, keep, r, scaling, p, w_is_KN, workspace, need_dx=True, dA=None, dB=None,
               accumulate=False):
    """dX (or None), dA [r,K] fp32, dB [N,r] fp32.  x2d is the forward's input, `packed` the forward's block."""
    _need_cuda(dy2d, x2d, W, packed, H, keep, workspace, dA, dB)
    M, K = x2d.shape
    N = dy2d.shape[1]
    dX = torch.empty_like(x2d) if need_dx else None
    if dA is None:
        dA = torch.empty((r, K), dtype=torch.float32, device=x2d.device)
    if dB is None:
        dB = torch.empty((N, r), dtype=torch.float32, device=x2d.device)
    rc = lib().lora_linear_bwd(
        _ptr(dy2d), _ptr(x2d), _ptr(W), _ptr(packed), _ptr(H), _ptr(keep), _ptr(dX), _ptr(dA), _ptr(dB), M, K, N,
        int(r), float(scaling), float(p), dtype_code(x2d.dtype), int(bool(w_is_KN)), int(bool(accumulate)),
        _ptr(workspace), workspace.numel() * workspace.element_size(), _stream(),
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
