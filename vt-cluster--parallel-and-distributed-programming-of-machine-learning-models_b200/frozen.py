"""The callers either side of the LoRA path, on the same kernels.

* `FrozenLinear(base, act=...)` with an activation: the frozen `mlp.c_fc` + gelu_new of a GPT-2 block
  (c_fc is not a LoRA target, finetune_lora_distilgpt2.py:100) or OPT's fc1 + relu when fc1 is not
  targeted (BASELINE configs[3]); the activation runs in the GEMM epilogue, its backward is one
  elementwise kernel.  `accelerate_frozen_neighbours(model)` does the module surgery.
* `FrozenLinear`: an un-adapted, frozen nn.Linear (in practice `lm_head`, which the reference leaves
  untouched: target_modules at finetune_lora_distilgpt2.py:100 / finetune_lora_opt_pp.py:109 do not
  name it) run through the tcgen05 GEMM of the LoRA path with the rank-r block off.  Its output may be
  ragged (vocab 50257): it is written into a buffer whose row stride is padded to 8 elements and
  returned as a strided view, so no copy and no mis-aligned library kernel.  Backward yields dX only.
* `fused_causal_lm_loss`: mean next-token cross-entropy over labels != -100, computed from the 16-bit
  logits in place of `logits.float()` + `F.cross_entropy` (DataCollatorForLanguageModeling semantics,
  finetune_lora_distilgpt2.py:114; shifted CE of finetune_lora_opt_pp.py:143-153).  The shift is
  applied to the labels, never to the logits.
No CPU fallback: both raise on CPU tensors / fp32.
"""
from __future__ import annotations

import ctypes

import torch
import torch.nn as nn

from . import binding


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(binding._stream())


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


def _on_device_of(argpos: int):
    """Decorator for autograd-node bodies: run with the device of positional argument `argpos` current, so that the
    kernels launch on that device's stream (a model on cuda:1 while cuda:0 is current must not launch on cuda:0)."""
    import functools

    def deco(fn):
        @functools.wraps(fn)
        def wrapped(*args, **kw):
            t = args[argpos]
            if not isinstance(t, torch.Tensor) or not t.is_cuda:
                raise RuntimeError("FrozenLinear / fused loss need CUDA tensors (no CPU fallback)")
            with binding.device_guard(t):
                return fn(*args, **kw)
        return wrapped
    return deco


class _FrozenLinearFn(torch.autograd.Function):
    @staticmethod
    @_on_device_of(1)
    def forward(ctx, x, W, bias, w_is_KN=0):
        if not x.is_cuda:
            raise RuntimeError("FrozenLinear needs CUDA tensors (no CPU fallback)")
        K = x.shape[-1]
        N = W.shape[1] if w_is_KN else W.shape[0]
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M = x2.shape[0]
        ldy = _pad8(N)
        ybuf = torch.empty((M, ldy), dtype=x.dtype, device=x.device)
        rc = binding.lib().lora_frozen_linear_fwd(_p(x2), K, _p(W), _p(bias), _p(ybuf), ldy, M, K, N,
                                                  binding.dtype_code(x.dtype), w_is_KN, _stream())
        binding._check(rc, "lora_frozen_linear_fwd")
        ctx.save_for_backward(W)
        ctx.meta = (x.shape, M, K, N, w_is_KN)
        return ybuf.view(*x.shape[:-1], ldy)[..., :N]  # strided view when N is ragged (row stride ldy)

    @staticmethod
    @_on_device_of(1)
    def backward(ctx, dy):
        (W,) = ctx.saved_tensors
        x_shape, M, K, N, w_is_KN = ctx.meta
        if not ctx.needs_input_grad[0]:
            return None, None, None, None
        try:
            dy2 = dy.view(M, N)  # keeps a padded-row layout (what the fused loss hands back) as it is
        except RuntimeError:
            dy2 = dy.reshape(M, N)
        if dy2.stride(1) != 1 or dy2.stride(0) % 8 != 0 or dy2.data_ptr() % 16 != 0:
            buf = torch.empty((M, _pad8(N)), dtype=dy.dtype, device=dy.device)
            buf[:, :N].copy_(dy2)
            dy2 = buf[:, :N]
        dX = torch.empty((M, K), dtype=dy.dtype, device=dy.device)
        rc = binding.lib().lora_frozen_linear_bwd(_p(dy2), dy2.stride(0), _p(W), _p(dX), K, M, K, N,
                                                  binding.dtype_code(dy.dtype), w_is_KN, _stream())
        binding._check(rc, "lora_frozen_linear_bwd")
        return dX.view(x_shape), None, None, None


class _FrozenLinearAddFn(torch.autograd.Function):
    """y = x W^T + bias + residual with the add in the GEMM epilogue (lora_frozen_linear_add_fwd); dX through
    lora_frozen_linear_bwd, the residual's gradient is dy itself."""

    @staticmethod
    @_on_device_of(1)
    def forward(ctx, x, W, bias, residual, w_is_KN=0):
        K = x.shape[-1]
        N = W.shape[1] if w_is_KN else W.shape[0]
        if N % 8 != 0:
            raise ValueError("frozen_linear_add: out_features must be a multiple of 8")
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M = x2.shape[0]
        r2 = residual.reshape(M, N)
        if not r2.is_contiguous():
            r2 = r2.contiguous()
        if r2.dtype != x.dtype or r2.device != x.device:
            raise ValueError("frozen_linear_add: residual must have the activation's dtype and device")
        y = torch.empty((M, N), dtype=x.dtype, device=x.device)
        rc = binding.lib().lora_frozen_linear_add_fwd(_p(x2), K, _p(W), _p(bias), _p(r2), N, _p(y), N, M, K, N,
                                                      binding.dtype_code(x.dtype), w_is_KN, _stream())
        binding._check(rc, "lora_frozen_linear_add_fwd")
        ctx.save_for_backward(W)
        ctx.meta = (x.shape, residual.shape, M, K, N, w_is_KN)
        return y.view(*x.shape[:-1], N)

    @staticmethod
    @_on_device_of(1)
    def backward(ctx, dy):
        (W,) = ctx.saved_tensors
        x_shape, r_shape, M, K, N, w_is_KN = ctx.meta
        dy2 = dy.reshape(M, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dX = None
        if ctx.needs_input_grad[0]:
            dX = torch.empty((M, K), dtype=dy.dtype, device=dy.device)
            rc = binding.lib().lora_frozen_linear_bwd(_p(dy2), N, _p(W), _p(dX), K, M, K, N,
                                                      binding.dtype_code(dy.dtype), w_is_KN, _stream())
            binding._check(rc, "lora_frozen_linear_bwd")
            dX = dX.view(x_shape)
        dR = dy2.view(r_shape) if ctx.needs_input_grad[3] else None
        return dX, None, None, dR, None


def frozen_linear_add(x, weight, bias, residual, w_is_KN: int = 0):
    """residual + linear(x) for a frozen projection, the add done in the GEMM epilogue.  What
    GPT2Block / OPTDecoderLayer compute as `residual + proj(h)` when no dropout sits between the two (eval, or
    dropout = 0; with dropout > 0 the HF blocks drop the projection's output first, and the fusion does not apply)."""
    return _FrozenLinearAddFn.apply(x, weight, bias, residual, w_is_KN)


ACT_CODES = {None: 0, "relu": 1, "gelu_new": 2}


def _weight_layout(base):
    """(in_features, out_features, w_is_KN) of nn.Linear ([N,K]) or HF Conv1D ([K,N])."""
    if isinstance(base, nn.Linear):
        return base.in_features, base.out_features, 0
    if type(base).__name__ == "Conv1D" and hasattr(base, "nf"):
        return base.weight.shape[0], base.weight.shape[1], 1
    raise ValueError(f"FrozenLinear wraps nn.Linear or transformers Conv1D, not {type(base).__name__}")


class _FrozenActLinearFn(torch.autograd.Function):
    """y = act(x W^T + b) for a frozen W; saves the pre-activation (gelu_new) or y (relu) for dX."""

    @staticmethod
    @_on_device_of(1)
    def forward(ctx, x, W, bias, act, w_is_KN):
        if not x.is_cuda:
            raise RuntimeError("FrozenLinear needs CUDA tensors (no CPU fallback)")
        K = x.shape[-1]
        N = W.shape[1] if w_is_KN else W.shape[0]
        if N % 8:
            raise RuntimeError("FrozenLinear with an activation needs out_features % 8 == 0")
        x2 = x.reshape(-1, K)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        M = x2.shape[0]
        need_grad = any(ctx.needs_input_grad)
        y = torch.empty((M, N), dtype=x.dtype, device=x.device)
        z = torch.empty((M, N), dtype=x.dtype, device=x.device) if (need_grad and act == 2) else None
        rc = binding.lib().lora_frozen_linear_act_fwd(_p(x2), K, _p(W), _p(bias), _p(y), N, _p(z), N, M, K, N, act,
                                                      binding.dtype_code(x.dtype), w_is_KN, _stream())
        binding._check(rc, "lora_frozen_linear_act_fwd")
        ctx.save_for_backward(W, z if act == 2 else y)
        ctx.meta = (x.shape, M, K, N, act, w_is_KN)
        return y.view(*x.shape[:-1], N)

    @staticmethod
    @_on_device_of(1)
    def backward(ctx, dy):
        W, saved = ctx.saved_tensors
        x_shape, M, K, N, act, w_is_KN = ctx.meta
        if not ctx.needs_input_grad[0]:
            return None, None, None, None, None
        dy2 = dy.reshape(M, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        code = binding.dtype_code(dy.dtype)
        dz = torch.empty((M, N), dtype=dy.dtype, device=dy.device)
        rc = binding.lib().lora_act_bwd(_p(dy2), _p(saved), _p(dz), M * N, act, code, _stream())
        binding._check(rc, "lora_act_bwd")
        dX = torch.empty((M, K), dtype=dy.dtype, device=dy.device)
        rc = binding.lib().lora_frozen_linear_bwd(_p(dz), N, _p(W), _p(dX), K, M, K, N, code, w_is_KN, _stream())
        binding._check(rc, "lora_frozen_linear_bwd")
        return dX.view(x_shape), None, None, None, None


class FrozenLinear(nn.Module):
    """Wraps a frozen nn.Linear or HF Conv1D (shares its Parameters, so weight tying survives);
    `act` in (None, "relu", "gelu_new") is applied in the GEMM epilogue."""

    def __init__(self, base: nn.Module, act=None):
        super().__init__()
        self.in_features, self.out_features, self.w_is_KN = _weight_layout(base)
        if base.weight.requires_grad or (base.bias is not None and base.bias.requires_grad):
            raise ValueError("FrozenLinear is for frozen layers: it produces no weight gradient")
        if act not in ACT_CODES:
            raise ValueError(f"act must be one of {list(ACT_CODES)}")
        self.base_layer = base
        self.act = act

    @property
    def weight(self):
        return self.base_layer.weight

    @property
    def bias(self):
        return self.base_layer.bias

    def forward(self, x):
        W = self.base_layer.weight
        if x.dtype != W.dtype:
            x = x.to(W.dtype)
        if self.act is None:
            return _FrozenLinearFn.apply(x, W, self.base_layer.bias, self.w_is_KN)
        return _FrozenActLinearFn.apply(x, W, self.base_layer.bias, ACT_CODES[self.act], self.w_is_KN)

    def extra_repr(self):
        return f"in_features={self.in_features}, out_features={self.out_features}, act={self.act}"


def _is_frozen_plain(m):
    try:
        _weight_layout(m)
    except ValueError:
        return False
    return not m.weight.requires_grad and (m.bias is None or not m.bias.requires_grad)


def accelerate_frozen_neighbours(model: nn.Module, plain: bool = False) -> int:
    """Module surgery on the un-adapted neighbours of the LoRA path; returns how many layers changed.
    GPT-2 blocks: frozen `mlp.c_fc` followed by NewGELUActivation -> one GEMM with gelu_new epilogue.
    OPT layers: frozen `fc1` followed by ReLU -> one GEMM with relu epilogue.
    plain=True also moves every other frozen nn.Linear / Conv1D inside those blocks onto the kernel."""
    n = 0
    for mod in list(model.modules()):
        kids = mod._modules
        if "c_fc" in kids and "act" in kids and _is_frozen_plain(kids["c_fc"]) and \
                type(kids["act"]).__name__ == "NewGELUActivation":
            mod.c_fc = FrozenLinear(kids["c_fc"], act="gelu_new")
            mod.act = nn.Identity()
            n += 1
        if "fc1" in kids and "activation_fn" in kids and _is_frozen_plain(kids["fc1"]) and \
                isinstance(kids["activation_fn"], nn.ReLU):
            mod.fc1 = FrozenLinear(kids["fc1"], act="relu")
            mod.activation_fn = nn.Identity()
            n += 1
        if plain and type(mod).__name__ in ("GPT2MLP", "GPT2Attention", "OPTAttention", "OPTDecoderLayer"):
            for name, kid in list(kids.items()):
                if _is_frozen_plain(kid) and kid.weight.shape[0] % 8 == 0 and kid.weight.shape[1] % 8 == 0:
                    setattr(mod, name, FrozenLinear(kid))
                    n += 1
    return n


def accelerate_lm_head(model: nn.Module) -> nn.Module:
    """Swap `lm_head` (found on the model or below a PeftModel wrapper) for a FrozenLinear."""
    for holder in (model, getattr(model, "base_model", None), getattr(getattr(model, "base_model", None), "model", None)):
        if holder is None:
            continue
        head = holder.__dict__.get("_modules", {}).get("lm_head")
        if isinstance(head, nn.Linear):
            holder.lm_head = FrozenLinear(head)
            return model
    raise ValueError("no nn.Linear lm_head found")


class _FusedCEFn(torch.autograd.Function):
    """sum_i loss_i / n_valid over rows of a [M, V] 16-bit logits matrix (row stride % 8 == 0)."""

    @staticmethod
    @_on_device_of(1)
    def forward(ctx, logits2, labels1, ignore_index):
        M, V = logits2.shape
        ld = logits2.stride(0)
        lse = torch.empty(M, dtype=torch.float32, device=logits2.device)
        rows = torch.empty(M, dtype=torch.float32, device=logits2.device)
        rc = binding.lib().lora_ce_fwd(_p(logits2), ld, _p(labels1), _p(lse), _p(rows), M, V, ignore_index,
                                       binding.dtype_code(logits2.dtype), _stream())
        binding._check(rc, "lora_ce_fwd")
        n_valid = (labels1 != ignore_index).sum().clamp_min(1).to(torch.float32)
        ctx.save_for_backward(logits2, labels1, lse, n_valid)
        ctx.ignore_index = ignore_index
        return rows.sum() / n_valid

    @staticmethod
    @_on_device_of(1)
    def backward(ctx, g):
        logits2, labels1, lse, n_valid = ctx.saved_tensors
        M, V = logits2.shape
        ldd = _pad8(V)
        dbuf = torch.empty((M, ldd), dtype=logits2.dtype, device=logits2.device)
        scale = (g.to(torch.float32) / n_valid).reshape(1).contiguous()
        rc = binding.lib().lora_ce_bwd(_p(logits2), logits2.stride(0), _p(labels1), _p(lse), _p(scale), _p(dbuf), ldd, M, V,
                                       ctx.ignore_index, binding.dtype_code(logits2.dtype), _stream())
        binding._check(rc, "lora_ce_bwd")
        return dbuf[:, :V], None, None


def fused_causal_lm_loss(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """logits [B, S, V] (16-bit, CUDA; may be a padded-row view), labels [B, S] -> scalar fp32 loss."""
    if not logits.is_cuda or logits.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("fused_causal_lm_loss needs 16-bit CUDA logits (no CPU fallback)")
    B, S, V = logits.shape
    shifted = torch.full((B, S), ignore_index, dtype=torch.long, device=logits.device)
    shifted[:, :-1] = labels[:, 1:]
    if logits.stride(-1) == 1 and logits.stride(1) % 8 == 0 and logits.stride(0) == S * logits.stride(1) and \
            logits.data_ptr() % 16 == 0:
        l2 = logits.view(B * S, V)  # rows stay where they are (padded row stride allowed)
    else:
        buf = torch.empty((B * S, _pad8(V)), dtype=logits.dtype, device=logits.device)
        buf[:, :V].copy_(logits.reshape(B * S, V))
        l2 = buf[:, :V]
    return _FusedCEFn.apply(l2, shifted.reshape(-1), int(ignore_index))


# ----------------------------------------------------------------------------------------------
# lm_head fused with the loss in token chunks (SURVEY.md 8f row N2 as written): the [M, V] logits never exist.
# ----------------------------------------------------------------------------------------------
class _ChunkedHeadCEFn(torch.autograd.Function):
    """loss = mean_{labels != ignore} CE(hidden W^T (+ b), shifted labels), computed chunk by chunk: for every
    block of `chunk` tokens the head GEMM writes its logits into ONE reused buffer (sized to stay in L2), the
    cross-entropy forward reads them, its backward overwrites them in place with dlogits / n_valid, and the
    head's dX GEMM consumes those -- four launches per chunk on the kernels above, no [M, V] tensor in HBM
    (distilgpt2 at 16 x 1024 tokens: 1.65 GB of logits and as much again of dlogits).  The weight is frozen, so the
    only gradient is dX; it is complete when forward returns, backward just scales it."""

    @staticmethod
    @_on_device_of(1)
    def forward(ctx, hidden, W, bias, labels1, ignore_index, chunk):
        K = hidden.shape[-1]
        V = W.shape[0]
        h2 = hidden.reshape(-1, K)
        if not h2.is_contiguous():
            h2 = h2.contiguous()
        M = h2.shape[0]
        dev, code = h2.device, binding.dtype_code(h2.dtype)
        ldv = _pad8(V)
        need_dx = ctx.needs_input_grad[0]
        lse = torch.empty(M, dtype=torch.float32, device=dev)
        rows = torch.empty(M, dtype=torch.float32, device=dev)
        n_valid = (labels1 != ignore_index).sum().clamp_min(1).to(torch.float32)
        scale = (1.0 / n_valid).reshape(1).contiguous()
        dX = torch.empty((M, K), dtype=h2.dtype, device=dev) if need_dx else None
        buf = torch.empty((min(chunk, M), ldv), dtype=h2.dtype, device=dev)
        lib, st, esz = binding.lib(), _stream(), h2.element_size()
        for c0 in range(0, M, chunk):
            m = min(chunk, M - c0)
            hp = ctypes.c_void_p(h2.data_ptr() + c0 * K * esz)
            lp = ctypes.c_void_p(labels1.data_ptr() + c0 * 8)
            lsep = ctypes.c_void_p(lse.data_ptr() + c0 * 4)
            rc = lib.lora_frozen_linear_fwd(hp, K, _p(W), _p(bias), _p(buf), ldv, m, K, V, code, 0, st)
            binding._check(rc, "lora_frozen_linear_fwd")
            rc = lib.lora_ce_fwd(_p(buf), ldv, lp, lsep, ctypes.c_void_p(rows.data_ptr() + c0 * 4), m, V, ignore_index, code, st)
            binding._check(rc, "lora_ce_fwd")
            if need_dx:
                rc = lib.lora_ce_bwd(_p(buf), ldv, lp, lsep, _p(scale), _p(buf), ldv, m, V, ignore_index, code, st)
                binding._check(rc, "lora_ce_bwd")
                rc = lib.lora_frozen_linear_bwd(_p(buf), ldv, _p(W), ctypes.c_void_p(dX.data_ptr() + c0 * K * esz), K, m, K, V,
                                                code, 0, st)
                binding._check(rc, "lora_frozen_linear_bwd")
        ctx.save_for_backward(dX)
        ctx.hshape = hidden.shape
        return rows.sum() / n_valid

    @staticmethod
    def backward(ctx, g):
        (dX,) = ctx.saved_tensors
        if dX is None:
            return None, None, None, None, None, None
        return (dX * g.to(dX.dtype)).view(ctx.hshape), None, None, None, None, None


def chunked_head_loss(hidden: torch.Tensor, weight: torch.Tensor, labels: torch.Tensor, bias=None,
                      ignore_index: int = -100, chunk_tokens: int = 0) -> torch.Tensor:
    """hidden [B, S, K] (16-bit, CUDA), weight [V, K] frozen, labels [B, S] -> scalar fp32 causal-LM loss
    (next-token, labels shifted here) without materialising the logits.  chunk_tokens = 0: as many tokens per
    chunk as keep the logits buffer within 96 MiB (it lives in the 126 MB L2 between the four kernels)."""
    if not hidden.is_cuda or hidden.dtype not in (torch.bfloat16, torch.float16):
        raise RuntimeError("chunked_head_loss needs 16-bit CUDA hidden states (no CPU fallback)")
    if weight.requires_grad:
        raise RuntimeError("chunked_head_loss is for a frozen head: it produces no weight gradient")
    B, S, _ = hidden.shape
    V = weight.shape[0]
    if chunk_tokens <= 0:
        chunk_tokens = max(256, ((96 << 20) // (_pad8(V) * hidden.element_size())) // 256 * 256)
    shifted = torch.full((B, S), ignore_index, dtype=torch.long, device=hidden.device)
    shifted[:, :-1] = labels[:, 1:]
    return _ChunkedHeadCEFn.apply(hidden, weight, bias, shifted.reshape(-1), int(ignore_index), int(chunk_tokens))


class DeferredHead(nn.Module):
    """Stands in for `lm_head` while training with `chunked_head_loss`: forward hands the hidden states through
    unchanged (the model's "logits" output then carries hidden states, which the paired loss function turns into the
    loss); in eval mode it is the plain frozen projection.  Shares the base layer's Parameters (weight tying survives)."""

    def __init__(self, base: nn.Linear):
        super().__init__()
        if base.weight.requires_grad:
            raise ValueError("DeferredHead is for a frozen lm_head")
        self.base_layer = base
        self._plain = FrozenLinear(base)

    @property
    def weight(self):
        return self.base_layer.weight

    @property
    def bias(self):
        return self.base_layer.bias

    def forward(self, hidden):
        return hidden if self.training else self._plain(hidden)


def fuse_lm_head_with_loss(model: nn.Module, chunk_tokens: int = 0):
    """Swap `lm_head` for a DeferredHead and return the loss function to pair with it:
    `loss_fn(model(**batch).logits, labels)` -- same call shape as fused_causal_lm_loss -- computes the chunked
    head + cross-entropy.  Training mode only (eval returns real logits)."""
    for holder in (model, getattr(model, "base_model", None), getattr(getattr(model, "base_model", None), "model", None)):
        if holder is None:
            continue
        head = holder.__dict__.get("_modules", {}).get("lm_head")
        if isinstance(head, FrozenLinear):
            head = head.base_layer
        if isinstance(head, nn.Linear):
            dh = DeferredHead(head)
            holder.lm_head = dh

            def loss_fn(hidden, labels, ignore_index: int = -100):
                return chunked_head_loss(hidden, dh.weight, labels, dh.bias, ignore_index, chunk_tokens)

            return loss_fn
    raise ValueError("no nn.Linear lm_head found")
