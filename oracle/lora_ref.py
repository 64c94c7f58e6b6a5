"""TEST INFRASTRUCTURE ONLY -- CPU fp32 restatement of the reference's LoRA hot path.

PARITY UNPINNED: the reference holds no golden vectors, known-answer tests or
saved adapters for this path, and the module that
owns the arithmetic -- third-party `peft` (lora.Linear), unpinned: it is absent
from /root/reference/requirements.txt:1-58 -- cannot be installed here.  This
file restates peft's published algorithm on the reference's own compute engine
(torch CPU fp32; requirements.txt:50, finetune_lora_distilgpt2.py:89,139) and
anchors on the reference's call sites:

  * hyper-parameters / targets: finetune_lora_distilgpt2.py:98-103
    (r=8, lora_alpha=16, target_modules=["c_attn","c_proj"], lora_dropout=0.05,
    bias="none") and finetune_lora_opt_pp.py:104-111 (six nn.Linear targets);
  * wrapping contract: get_peft_model at finetune_lora_distilgpt2.py:104 and
    finetune_lora_opt_pp.py:112, `.base_model.model` at finetune_lora_opt_pp.py:116,
    `.lm_head` forwarding at finetune_lora_opt_pp.py:142;
  * base layers: transformers.pytorch_utils.Conv1D (weight [K,N], addmm) for
    GPT-2, nn.Linear (weight [N,K]) for OPT.

peft lora.Linear semantics restated (upstream peft/tuners/lora/layer.py):
    result = base_layer(x) + lora_B(lora_A(dropout(x))) * scaling
    scaling = lora_alpha / r;  lora_A.weight [r,K] ~ kaiming_uniform(a=sqrt(5));
    lora_B.weight [N,r] = 0;  every non-LoRA parameter frozen;  Conv1D targets
    set fan_in_fan_out=True (the base weight is stored transposed).
Gradients come from torch autograd over exactly those ops, as in the reference.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import
this module.  The product path never does.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Iterable, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F


# --------------------------------------------------------------------------
# functional form (used by the kernel parity tests and the microbench baseline)
# --------------------------------------------------------------------------
def lora_linear_forward(x, W, bias, A, B, scaling, keep=None, p=0.0, w_is_KN=False):
    """y = base(x) + scaling * ((keep * x / (1-p)) A^T) B^T, all fp32 torch CPU ops.

    x [M,K]; W [N,K] (nn.Linear) or [K,N] (Conv1D, w_is_KN); A [r,K]; B [N,r];
    keep: optional bool/0-1 tensor [M,K] (explicit dropout mask, shared with the GPU side).
    """
    if w_is_KN:
        base = torch.addmm(bias, x, W) if bias is not None else x @ W  # Conv1D.forward
    else:
        base = F.linear(x, W, bias)  # nn.Linear.forward
    xd = x
    if keep is not None and p > 0.0:
        xd = x * keep.to(x.dtype) / (1.0 - p)  # F.dropout with a given mask
    h = F.linear(xd, A)
    return base + F.linear(h, B) * scaling, h


def lora_linear_fwd_bwd(x, W, bias, A, B, scaling, dy, keep=None, p=0.0, w_is_KN=False):
    """Forward + autograd backward; returns dict(y, h, dx, dA, dB).  No dW/dbias: base is frozen."""
    x = x.detach().clone().requires_grad_(True)
    A = A.detach().clone().requires_grad_(True)
    B = B.detach().clone().requires_grad_(True)
    y, h = lora_linear_forward(x, W, bias, A, B, scaling, keep, p, w_is_KN)
    y.backward(dy)
    return {"y": y.detach(), "h": h.detach(), "dx": x.grad, "dA": A.grad, "dB": B.grad}


# --------------------------------------------------------------------------
# module form: a peft-free injector with peft's attribute / state-dict names
# --------------------------------------------------------------------------
@dataclass
class RefLoraConfig:
    r: int = 8
    lora_alpha: int = 8
    target_modules: Iterable[str] = field(default_factory=list)
    lora_dropout: float = 0.0
    bias: str = "none"
    task_type: Optional[str] = None
    fan_in_fan_out: bool = False


def _is_conv1d(m: nn.Module) -> bool:
    return type(m).__name__ == "Conv1D" and hasattr(m, "nf")


class RefLoraLinear(nn.Module):
    """peft lora.Linear restated for one adapter named "default"."""

    def __init__(self, base_layer: nn.Module, cfg: RefLoraConfig):
        super().__init__()
        self.base_layer = base_layer
        if _is_conv1d(base_layer):
            K, N = base_layer.weight.shape
            self.fan_in_fan_out = True
        else:
            N, K = base_layer.weight.shape
            self.fan_in_fan_out = False
        self.in_features, self.out_features = K, N
        self.r = {"default": cfg.r}
        self.lora_alpha = {"default": cfg.lora_alpha}
        self.scaling = {"default": cfg.lora_alpha / cfg.r}
        self.lora_dropout = nn.ModuleDict(
            {"default": nn.Dropout(cfg.lora_dropout) if cfg.lora_dropout > 0 else nn.Identity()}
        )
        self.lora_A = nn.ModuleDict({"default": nn.Linear(K, cfg.r, bias=False)})
        self.lora_B = nn.ModuleDict({"default": nn.Linear(cfg.r, N, bias=False)})
        nn.init.kaiming_uniform_(self.lora_A["default"].weight, a=math.sqrt(5))
        nn.init.zeros_(self.lora_B["default"].weight)
        self.p = float(cfg.lora_dropout)
        # Test hook: when set to a callable (M, K) -> bool tensor, the mask is taken from it
        # instead of torch's CPU generator, so that CPU and GPU runs share masks.
        self.mask_fn = None

    def forward(self, x):
        result = self.base_layer(x)
        A = self.lora_A["default"].weight
        B = self.lora_B["default"].weight
        if self.training and self.p > 0 and self.mask_fn is not None:
            x2 = x.reshape(-1, x.shape[-1])
            keep = self.mask_fn(x2.shape[0], x2.shape[1]).to(x.dtype)
            xd = (x2 * keep / (1.0 - self.p)).reshape(x.shape)
        else:
            xd = self.lora_dropout["default"](x)
        return result + F.linear(F.linear(xd, A), B) * self.scaling["default"]


class RefPeftModel(nn.Module):
    """Shape of peft's PeftModelForCausalLM that the reference relies on:
    `.base_model.model` is the original HF model (finetune_lora_opt_pp.py:116) and unknown
    attributes forward to it (`.lm_head`, finetune_lora_opt_pp.py:142)."""

    class _Tuner(nn.Module):
        def __init__(self, model):
            super().__init__()
            self.model = model

        def forward(self, *a, **kw):
            return self.model(*a, **kw)

    def __init__(self, model, cfg):
        super().__init__()
        self.base_model = RefPeftModel._Tuner(model)
        self.peft_config = {"default": cfg}

    def forward(self, *a, **kw):
        return self.base_model(*a, **kw)

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(self.base_model.model, name)


def ref_get_peft_model(model: nn.Module, cfg: RefLoraConfig) -> RefPeftModel:
    """Suffix-match target modules, wrap them, freeze everything that is not LoRA."""
    targets = list(cfg.target_modules)
    hits = [
        name
        for name, m in model.named_modules()
        if (isinstance(m, nn.Linear) or _is_conv1d(m)) and any(name == t or name.endswith("." + t) for t in targets)
    ]
    for name in hits:
        parent_name, _, child = name.rpartition(".")
        parent = model.get_submodule(parent_name) if parent_name else model
        setattr(parent, child, RefLoraLinear(getattr(parent, child), cfg))
    for n, p in model.named_parameters():
        p.requires_grad_("lora_" in n)
    return RefPeftModel(model, cfg)


# --------------------------------------------------------------------------
# Trainer-equivalent step loop (what HF Trainer does with the reference's
# TrainingArguments, finetune_lora_distilgpt2.py:130-145): AdamW lr 5e-5,
# betas (0.9, 0.999), eps 1e-8, weight_decay 0, linear decay to 0, grad-norm
# clip 1.0, gradient accumulation, loss = mean CE over shifted tokens != -100.
# --------------------------------------------------------------------------
def causal_lm_loss(logits, labels):
    shift_logits = logits[:, :-1, :].contiguous().float()
    shift_labels = labels[:, 1:].contiguous()
    return F.cross_entropy(shift_logits.view(-1, shift_logits.size(-1)), shift_labels.view(-1), ignore_index=-100)


def train_steps(model, batches, lr=5e-5, accum=1, total_steps=None, max_grad_norm=1.0, reduce_fn=None):
    """Run len(batches)//accum optimiser steps; returns the per-step mean losses.

    `batches` is a list of dicts(input_ids, attention_mask, labels); `reduce_fn(params)` (optional)
    averages gradients across ranks before clipping (the DDP all-reduce of finetune_lora_distilgpt2.py:139-144).
    """
    params = [p for p in model.parameters() if p.requires_grad]
    opt = torch.optim.AdamW(params, lr=lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0)
    n_steps = len(batches) // accum
    total = total_steps or n_steps
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: max(0.0, (total - s) / max(1, total)))
    losses = []
    model.train()
    for step in range(n_steps):
        acc = 0.0
        for micro in range(accum):
            b = batches[step * accum + micro]
            out = model(input_ids=b["input_ids"], attention_mask=b.get("attention_mask"))
            logits = out.logits if hasattr(out, "logits") else out
            loss = causal_lm_loss(logits, b["labels"]) / accum
            loss.backward()
            acc += float(loss.detach())
        if reduce_fn is not None:
            reduce_fn(params)
        torch.nn.utils.clip_grad_norm_(params, max_grad_norm)
        opt.step()
        sched.step()
        opt.zero_grad(set_to_none=True)
        losses.append(acc)
    return losses
