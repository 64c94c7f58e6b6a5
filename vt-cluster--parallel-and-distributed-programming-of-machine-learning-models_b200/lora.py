"""LoRA-adapted projection backed by the sm_100a kernels.

`Linear` mirrors peft's `lora.Linear` (third-party; reached by the reference
through get_peft_model at finetune_lora_distilgpt2.py:104 and
finetune_lora_opt_pp.py:112): same attribute names (`base_layer`, `lora_A`,
`lora_B`, `lora_dropout`, `scaling`, `r`, `lora_alpha`, `fan_in_fan_out`,
`merged`, `active_adapters`), same state-dict keys, same initialisation
(A kaiming-uniform(a=sqrt(5)), B zero), same forward semantics
    result = base_layer(x) + lora_B(lora_A(dropout(x))) * scaling.
The arithmetic runs in ONE autograd node over the C ABI in include/lora_b200.h;
there is no eager/PyTorch fallback.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
import torch.nn as nn

from . import binding

_WORKSPACES: dict = {}


def _workspace(device: torch.device, nbytes: int) -> torch.Tensor:
    """Scratch of the backward, one block per (device, stream), grown on demand: kernels of one stream serialise
    on it, backward passes running on different streams never share one."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    key = (idx, binding._stream(device))
    ws = _WORKSPACES.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.zeros(max(nbytes, 1 << 20), dtype=torch.uint8, device=device)
        _WORKSPACES[key] = ws
    return ws


def _philox_seed_offset(device: torch.device, numel: int):
    """Take (seed, offset) from torch's CUDA generator and advance it, the way ATen's own
    dropout kernels do.  torch.utils.checkpoint restores this state before recomputing, so the
    recomputed forward (gradient checkpointing is on in both reference scripts,
    finetune_lora_distilgpt2.py:93-96, finetune_lora_opt_pp.py:99-102) redraws the same mask."""
    idx = device.index if device.index is not None else torch.cuda.current_device()
    gen = torch.cuda.default_generators[idx]
    seed = gen.initial_seed()
    offset = gen.get_offset()
    gen.set_offset(offset + 4 * ((numel + 31) // 32))  # multiple of 4, as ATen requires
    return seed & 0xFFFFFFFFFFFFFFFF, offset


class _FusedLoraLinearFn(torch.autograd.Function):
    """y = x W^T + b + s (drop(x) A^T) B^T ; backward gives dx, dA, dB and nothing for W, b."""

    @staticmethod
    def forward(ctx, x, W, bias, A, B, scaling, p, w_is_KN, seed, offset):
        K = x.shape[-1]
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        N = W.shape[1] if w_is_KN else W.shape[0]
        with binding.device_guard(x2):
            Y, H, keep, packed = binding.linear_fwd(x2, W, bias, A, B, scaling, p, seed, offset, w_is_KN)
        # Saved for backward: the input itself (shared with every other projection that reads it -- q/k/v save ONE
        # x, where stock autograd saves a dropout(x) per adapter), H, the keep-bits (1/16 of x) and the 16-bit
        # adapter operands this forward packed (backward differentiates with the values the forward used).
        ctx.save_for_backward(x2, W, H, keep, packed)
        ctx.meta = (scaling, p, w_is_KN, x.shape, N, A.shape[0])
        return Y.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, W, H, keep, packed = ctx.saved_tensors
        scaling, p, w_is_KN, x_shape, N, r = ctx.meta
        dy2 = dy.reshape(-1, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        M, K = x2.shape
        need_dx = ctx.needs_input_grad[0]
        with binding.device_guard(x2):
            ws = _workspace(x2.device, binding.workspace_bytes(M, K, N, r))
            dX, dA, dB = binding.linear_bwd(dy2, x2, W, packed, H, keep, r, scaling, p, w_is_KN, ws, need_dx=need_dx)
        return (dX.view(x_shape) if need_dx else None, None, None, dA, dB, None, None, None, None, None)


def fused_lora_linear(x, W, bias, A, B, scaling, p=0.0, training=True, w_is_KN=False, seed=None, offset=None):
    """Functional entry: the call a user of this library makes for one adapted projection."""
    if x.dtype != W.dtype:
        x = x.to(W.dtype)
    if A.dtype != torch.float32:
        A = A.float()
    if B.dtype != torch.float32:
        B = B.float()
    p_eff = float(p) if training else 0.0
    if p_eff > 0.0 and seed is None:
        seed, offset = _philox_seed_offset(x.device, x.numel())
    return _FusedLoraLinearFn.apply(x, W, bias, A.contiguous(), B.contiguous(), float(scaling), p_eff, bool(w_is_KN),
                                    int(seed or 0), int(offset or 0))


def _is_conv1d(m: nn.Module) -> bool:
    return type(m).__name__ == "Conv1D" and hasattr(m, "nf")


class Linear(nn.Module):
    """Drop-in for peft.tuners.lora.Linear over an nn.Linear or a GPT-2 Conv1D base layer."""

    def __init__(self, base_layer: nn.Module, adapter_name: str = "default", r: int = 8, lora_alpha: int = 8,
                 lora_dropout: float = 0.0, fan_in_fan_out: bool = False, init_lora_weights: bool = True,
                 use_rslora: bool = False):
        super().__init__()
        if isinstance(base_layer, nn.Linear):
            out_features, in_features = base_layer.weight.shape
            if fan_in_fan_out:
                raise ValueError("fan_in_fan_out=True is for Conv1D base layers; nn.Linear stores W as [out, in]")
        elif _is_conv1d(base_layer):
            in_features, out_features = base_layer.weight.shape
            fan_in_fan_out = True  # peft flips this for Conv1D and warns
        else:
            raise ValueError(f"unsupported LoRA target {type(base_layer).__name__}: nn.Linear or Conv1D only")
        if not 1 <= r <= 64:
            raise ValueError(f"rank r must be in [1, 64], got {r}")
        self.base_layer = base_layer
        self.in_features, self.out_features = in_features, out_features
        self.fan_in_fan_out = fan_in_fan_out
        self.r, self.lora_alpha, self.scaling = {}, {}, {}
        self.lora_dropout = nn.ModuleDict({})
        self.lora_A = nn.ModuleDict({})
        self.lora_B = nn.ModuleDict({})
        self.merged_adapters: list = []
        self._disable_adapters = False
        self._active_adapter = adapter_name
        self.update_layer(adapter_name, r, lora_alpha, lora_dropout, init_lora_weights, use_rslora)

    # ---- peft-compatible surface -------------------------------------------------
    @property
    def weight(self):
        return self.base_layer.weight

    @property
    def bias(self):
        return self.base_layer.bias

    @property
    def merged(self) -> bool:
        return bool(self.merged_adapters)

    @property
    def active_adapters(self):
        return [self._active_adapter]

    @property
    def active_adapter(self):
        return self._active_adapter

    @property
    def disable_adapters(self) -> bool:
        return self._disable_adapters

    def enable_adapters(self, enabled: bool):
        self._disable_adapters = not enabled
        for d in (self.lora_A, self.lora_B):
            for m in d.values():
                m.requires_grad_(enabled)

    def update_layer(self, adapter_name, r, lora_alpha, lora_dropout, init_lora_weights=True, use_rslora=False):
        self.r[adapter_name] = r
        self.lora_alpha[adapter_name] = lora_alpha
        self.lora_dropout[adapter_name] = nn.Dropout(p=lora_dropout) if lora_dropout > 0.0 else nn.Identity()
        dev = self.base_layer.weight.device
        # Adapter masters stay fp32 whatever the base dtype (peft's autocast_adapter_dtype default).
        self.lora_A[adapter_name] = nn.Linear(self.in_features, r, bias=False, device=dev, dtype=torch.float32)
        self.lora_B[adapter_name] = nn.Linear(r, self.out_features, bias=False, device=dev, dtype=torch.float32)
        self.scaling[adapter_name] = lora_alpha / math.sqrt(r) if use_rslora else lora_alpha / r
        if init_lora_weights:
            nn.init.kaiming_uniform_(self.lora_A[adapter_name].weight, a=math.sqrt(5))
            nn.init.zeros_(self.lora_B[adapter_name].weight)

    def get_delta_weight(self, adapter: str) -> torch.Tensor:
        """s * B A in the base weight's layout and dtype (set-up time only; plain torch)."""
        A = self.lora_A[adapter].weight.float()
        B = self.lora_B[adapter].weight.float()
        delta = (B @ A) * self.scaling[adapter]
        if self.fan_in_fan_out:
            delta = delta.t()
        return delta.to(self.base_layer.weight.dtype)

    def merge(self, safe_merge: bool = False, adapter_names=None):
        for name in adapter_names or self.active_adapters:
            if name in self.merged_adapters:
                continue
            if safe_merge and not torch.isfinite(self.get_delta_weight(name)).all():
                raise ValueError(f"NaNs detected in the merged weights of adapter {name}")
            self._add_delta(name, +1.0)
            self.merged_adapters.append(name)

    def unmerge(self):
        while self.merged_adapters:
            self._add_delta(self.merged_adapters.pop(), -1.0)

    def _add_delta(self, name: str, sign: float):
        """W += sign * s * B A.  16-bit CUDA weights: the in-place kernel (fp32 product, one rounding);
        anything else (fp32 / CPU set-up code and the CPU unit tests): plain torch."""
        W = self.base_layer.weight
        A, B = self.lora_A[name].weight, self.lora_B[name].weight
        if W.is_cuda and W.dtype in (torch.bfloat16, torch.float16) and W.is_contiguous():
            import ctypes

            A32, B32 = A.detach().float().contiguous(), B.detach().float().contiguous()
            with binding.device_guard(W):
                rc = binding.lib().lora_merge_weight(
                    ctypes.c_void_p(W.data_ptr()), ctypes.c_void_p(A32.data_ptr()), ctypes.c_void_p(B32.data_ptr()),
                    self.in_features, self.out_features, A.shape[0], float(sign * self.scaling[name]),
                    binding.dtype_code(W.dtype), int(bool(self.fan_in_fan_out)), ctypes.c_void_p(binding._stream(W.device)))
            binding._check(rc, "lora_merge_weight")
        else:
            W.data += sign * self.get_delta_weight(name)

    # ---- forward -------------------------------------------------------------------
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        W = self.base_layer.weight
        b = self.base_layer.bias
        if self._disable_adapters or self.merged:
            if self._disable_adapters and self.merged:
                self.unmerge()
            if torch.is_grad_enabled() and x.requires_grad:
                # gradient w.r.t. the input only (peft: the frozen base layer alone), same tcgen05 GEMM both ways
                from .frozen import _FrozenLinearFn

                return _FrozenLinearFn.apply(x.to(W.dtype), W, b, int(bool(self.fan_in_fan_out)))
            x2 = x.reshape(-1, x.shape[-1]).to(W.dtype).contiguous()
            with binding.device_guard(x2):
                y = binding.linear_base(x2, W, b, self.fan_in_fan_out)
            return y.view(*x.shape[:-1], self.out_features)
        name = self._active_adapter
        drop = self.lora_dropout[name]
        p = drop.p if isinstance(drop, nn.Dropout) else 0.0
        return fused_lora_linear(x, W, b, self.lora_A[name].weight, self.lora_B[name].weight, self.scaling[name], p=p,
                                 training=self.training, w_is_KN=self.fan_in_fan_out)

    def __repr__(self) -> str:
        return "lora." + super().__repr__()
