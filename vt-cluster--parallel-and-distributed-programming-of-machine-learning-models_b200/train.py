"""Step loop and data-parallel wrapper around a LoRA-adapted causal LM.

Restates, for a single 8xB200 box, what the reference gets from HF Trainer + accelerate with
its TrainingArguments (finetune_lora_distilgpt2.py:130-158): AdamW(lr, betas (0.9, 0.999),
eps 1e-8, weight_decay 0), linear LR decay to zero, gradient-norm clip 1.0, gradient
accumulation with the all-reduce skipped on non-boundary micro-steps, loss = mean
cross-entropy over shifted tokens whose label is not -100 (DataCollatorForLanguageModeling
with mlm=False, finetune_lora_distilgpt2.py:114).  Under DDP only the adapter parameters
(requires_grad) enter the reducer; the bucketed all-reduce runs on NCCL over NVLink and
overlaps the remaining backward (torch DDP's autograd hooks fire as each fused LoRA node
returns its dA/dB).

This module is host-side plumbing: it runs on whatever device the model lives on, so the CPU
test-suite can drive it with gloo (world_size 2) and the oracle's layer.
"""
from __future__ import annotations

import contextlib
import os
from dataclasses import dataclass
from typing import Callable, Dict, Iterable, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class TrainConfig:
    lr: float = 5e-5  # finetune_lora_distilgpt2.py:16
    accum: int = 32  # finetune_lora_distilgpt2.py:15 (--accum)
    max_grad_norm: float = 1.0  # HF Trainer default; DeepSpeed gradient_clipping 1.0 (finetune_lora_opt_pp.py:197)
    total_steps: Optional[int] = None  # for the linear decay; None -> number of steps run
    betas: tuple = (0.9, 0.999)
    eps: float = 1e-8
    weight_decay: float = 0.0
    constant_lr: bool = False  # the OPT pipeline script keeps lr constant (finetune_lora_opt_pp.py:207)
    fused_optimizer: bool = True  # CUDA: clip + AdamW as two launches of this library; False: torch.optim.AdamW


def causal_lm_loss(logits: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    shift_logits = logits[:, :-1, :].contiguous().float()
    shift_labels = labels[:, 1:].contiguous()
    return F.cross_entropy(shift_logits.view(-1, shift_logits.size(-1)), shift_labels.view(-1), ignore_index=-100)


def synthetic_batches(n: int, batch: int, seq: int, vocab: int, pad_id: int, seed: int, device="cpu") -> List[Dict]:
    """Random token batches of the named shape (no network for datasets): ids exclude the pad id,
    attention mask all ones, labels = ids (nothing padded, so nothing is -100)."""
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        ids = torch.randint(0, vocab - 1, (batch, seq), generator=g)
        ids = ids + (ids >= pad_id).long()  # skip pad_id
        out.append({"input_ids": ids.to(device), "attention_mask": torch.ones_like(ids).to(device),
                    "labels": ids.clone().to(device)})
    return out


def trainable_parameters(model: nn.Module):
    return [p for p in model.parameters() if p.requires_grad]


def wrap_ddp(model: nn.Module, device: Optional[torch.device] = None, bucket_cap_mb: int = 25,
             static_graph: bool = False, exchange: str = "nccl") -> nn.Module:
    """torch DDP the way HF Trainer configures it for the reference
    (ddp_find_unused_parameters=False, finetune_lora_distilgpt2.py:142): frozen base weights
    never enter the reducer, so the only traffic is the adapter-gradient bucket
    (1.6 MB for distilgpt2 r=8).  exchange="p2p" (CUDA only) replaces the NCCL all-reduce of each
    bucket by the in-stream peer-memory kernel of comm.PeerBucketExchange."""
    from torch.nn.parallel import DistributedDataParallel as DDP

    kw = dict(find_unused_parameters=False, bucket_cap_mb=bucket_cap_mb, gradient_as_bucket_view=True,
              static_graph=static_graph)
    if exchange not in ("nccl", "p2p"):
        raise ValueError("exchange must be 'nccl' or 'p2p'")
    if device is not None and device.type == "cuda":
        ddp = DDP(model, device_ids=[device.index], output_device=device.index, **kw)
        if exchange == "p2p":
            from .comm import try_peer_exchange

            n = sum(p.numel() for p in model.parameters() if p.requires_grad)
            ddp._lora_exchange = try_peer_exchange(n)  # kept alive with the wrapper; None: NCCL stays
            if ddp._lora_exchange is not None:
                ddp.register_comm_hook(None, ddp._lora_exchange.ddp_hook)
        return ddp
    if exchange == "p2p":
        raise RuntimeError("exchange='p2p' needs CUDA devices (no CPU fallback)")
    return DDP(model, **kw)


def close_exchange(ddp: nn.Module) -> None:
    """Collective: release the peer-memory exchange a `wrap_ddp(..., exchange="p2p")` wrapper holds (its own
    buffer and the IPC mappings of every peer's).  Call it on every rank before dropping the wrapper; a wrapper
    that used NCCL has nothing to release."""
    ex = getattr(ddp, "_lora_exchange", None)
    if ex is not None:
        ex.close()
        ddp._lora_exchange = None


class FusedAdamW:
    """Gradient-norm clip + AdamW of every adapter tensor in two kernel launches (C ABI lora_adamw_clip_step,
    csrc/optim.cuh) instead of the ~30 small kernels of clip_grad_norm_ + torch.optim.AdamW: on distilgpt2 the
    optimiser step is otherwise launch-bound.  Same formulas and defaults as the torch pair (HF Trainer's, which the
    reference uses: finetune_lora_distilgpt2.py:130-145).  CUDA fp32 parameters only; no CPU path."""

    def __init__(self, params, lr=5e-5, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_grad_norm=1.0):
        from . import binding

        self.params = [p for p in params]
        if not self.params or any((not p.is_cuda) or p.dtype != torch.float32 or not p.is_contiguous() for p in self.params):
            raise RuntimeError("FusedAdamW needs contiguous fp32 CUDA parameters (no CPU fallback)")
        self.lr, self.betas, self.eps, self.weight_decay, self.max_grad_norm = lr, betas, eps, weight_decay, max_grad_norm
        self.device = self.params[0].device
        self.exp_avg = [torch.zeros_like(p) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p) for p in self.params]
        self._steps = torch.zeros(len(self.params), dtype=torch.int32, device=self.device)  # per tensor, as torch keeps them
        self._binding = binding
        self._key, self._table, self._n_chunks, self._scratch = None, None, 0, None

    def _build_table(self, live):
        chunk = 4096  # LORA_ADAMW_CHUNK
        rows, first = [], 0
        for i in live:
            p = self.params[i]
            rows.append([p.data_ptr(), p.grad.data_ptr(), self.exp_avg[i].data_ptr(), self.exp_avg_sq[i].data_ptr(),
                         p.numel(), first, self._steps.data_ptr() + 4 * i])
            first += (p.numel() + chunk - 1) // chunk
        self._table = torch.tensor(rows, dtype=torch.int64).to(self.device)
        if self._scratch is None or self._scratch.numel() < 4 + first:
            self._scratch = torch.zeros(4 + first, dtype=torch.float32, device=self.device)
        self._n_chunks = first

    def step(self, lr=None):
        """One optimiser step over the parameters that have a gradient; returns the pre-clip gradient norm (a CUDA
        scalar view, no host sync)."""
        import ctypes

        live = [i for i, p in enumerate(self.params) if p.grad is not None]
        if not live:
            return None
        for i in live:
            g = self.params[i].grad
            if g.dtype != torch.float32 or not g.is_contiguous():
                self.params[i].grad = g = g.float().contiguous()
        key = tuple((i, self.params[i].data_ptr(), self.params[i].grad.data_ptr()) for i in live)
        if key != self._key:
            self._build_table(live)
            self._key = key
        b = self._binding
        with b.device_guard(self._table):
            rc = b.lib().lora_adamw_clip_step(
                ctypes.c_void_p(self._table.data_ptr()), len(live), self._n_chunks, float(self.lr if lr is None else lr),
                float(self.betas[0]), float(self.betas[1]), float(self.eps), float(self.weight_decay),
                float(self.max_grad_norm or 0.0), ctypes.c_void_p(self._scratch.data_ptr()),
                ctypes.c_void_p(b._stream(self.device)))
        b._check(rc, "lora_adamw_clip_step")
        return self._scratch[0]

    def zero_grad(self, set_to_none=True):
        for p in self.params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()


class Trainer:
    """Minimal stand-in for transformers.Trainer on this path (model(**batch) -> loss, backward,
    accumulate, reduce, clip, AdamW, linear decay)."""

    def __init__(self, model: nn.Module, cfg: TrainConfig, on_step_end: Optional[Callable] = None,
                 loss_fn: Optional[Callable] = None):
        self.loss_fn = loss_fn or causal_lm_loss  # e.g. frozen.fused_causal_lm_loss on the GPU
        self.model = model
        self.cfg = cfg
        self.params = trainable_parameters(model)
        # On CUDA the clip + AdamW pair runs as two launches of this library (FusedAdamW); elsewhere (the CPU tests,
        # the oracle runs) torch's own optimiser.
        self.fused = bool(cfg.fused_optimizer) and all(p.is_cuda and p.dtype == torch.float32 for p in self.params)
        if self.fused:
            self.opt = FusedAdamW(self.params, lr=cfg.lr, betas=cfg.betas, eps=cfg.eps, weight_decay=cfg.weight_decay,
                                  max_grad_norm=cfg.max_grad_norm)
        else:
            self.opt = torch.optim.AdamW(self.params, lr=cfg.lr, betas=cfg.betas, eps=cfg.eps,
                                         weight_decay=cfg.weight_decay)
        self._lr_fn = None
        self.sched = None
        self.global_step = 0
        self.on_step_end = on_step_end

    def _ensure_sched(self, n_steps: int):
        if self.sched is None:
            total = self.cfg.total_steps or n_steps
            if self.cfg.constant_lr:
                fn = lambda s: 1.0
            else:
                fn = lambda s: max(0.0, (total - s) / max(1, total))
            self._lr_fn = fn
            self.sched = True if self.fused else torch.optim.lr_scheduler.LambdaLR(self.opt, fn)

    def micro_step(self, batch: Dict, sync: bool) -> torch.Tensor:
        ctx = contextlib.nullcontext()
        if not sync and hasattr(self.model, "no_sync"):
            ctx = self.model.no_sync()  # accumulation micro-step: no all-reduce
        with ctx:
            out = self.model(input_ids=batch["input_ids"], attention_mask=batch.get("attention_mask"))
            logits = out.logits if hasattr(out, "logits") else out
            loss = self.loss_fn(logits, batch["labels"]) / self.cfg.accum
            loss.backward()
        return loss.detach()

    def optimizer_step(self):
        if self.fused:
            # LambdaLR semantics: the k-th step (k = 0, 1, ...) runs at lr * fn(k)
            self.last_grad_norm = self.opt.step(lr=self.cfg.lr * self._lr_fn(self.global_step))
        else:
            self.last_grad_norm = torch.nn.utils.clip_grad_norm_(self.params, self.cfg.max_grad_norm)
            self.opt.step()
            self.sched.step()
        self.opt.zero_grad(set_to_none=True)
        self.global_step += 1

    def train(self, batches: Iterable[Dict]) -> List[torch.Tensor]:
        batches = list(batches)
        accum = self.cfg.accum
        if len(batches) % accum != 0:
            # HF Trainer would still step on a short last accumulation group; this loop divides every loss by
            # `accum`, so a silent partial group would be mis-scaled.  Refuse instead of dropping batches.
            raise ValueError(f"{len(batches)} micro-batches is not a multiple of accum={accum}")
        n_steps = len(batches) // accum
        self._ensure_sched(n_steps)
        self.model.train()
        losses = []
        for step in range(n_steps):
            total = None
            for micro in range(accum):
                l = self.micro_step(batches[step * accum + micro], sync=(micro == accum - 1))
                total = l if total is None else total + l
            self.optimizer_step()
            losses.append(total)  # stays on the device: no host sync inside the loop
            if self.on_step_end is not None:
                self.on_step_end(self, step, total)
        return losses


def save_run(model: nn.Module, save_dir: str, meta: Optional[Dict] = None, tokenizer=None,
             meta_name: str = "run_meta.json") -> None:
    """Rank-0-only adapter save between two barriers, as finetune_lora_distilgpt2.py:162-189 does it:
    `save_pretrained` (adapter_config.json + adapter_model.safetensors), the tokenizer when given, and the
    run's metadata (base model, world size, seq_len, batch, accum, lr, train_seconds, ...) as JSON."""
    import json

    on = dist.is_available() and dist.is_initialized()
    if on:
        dist.barrier()
    close_exchange(model)
    if not on or dist.get_rank() == 0:
        core = model.module if hasattr(model, "module") and hasattr(model.module, "save_pretrained") else model
        core.save_pretrained(save_dir)
        if tokenizer is not None:
            tokenizer.save_pretrained(save_dir)
        if meta is not None:
            with open(os.path.join(save_dir, meta_name), "w") as f:
                json.dump(meta, f, indent=2)
    if on:
        dist.barrier()


def init_distributed(backend: Optional[str] = None):
    """env:// rendezvous as in finetune_lora_distilgpt2.py:35-45, NCCL on GPUs (gloo only in the
    CPU tests).  Returns (rank, world, local_rank)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kw = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend, init_method="env://", rank=rank, world_size=world, **kw)
        x = torch.ones(1, device=kw.get("device_id", "cpu"))
        dist.all_reduce(x)  # collective sanity check, finetune_lora_distilgpt2.py:50-54
        assert int(x.item()) == world
    return rank, world, local
