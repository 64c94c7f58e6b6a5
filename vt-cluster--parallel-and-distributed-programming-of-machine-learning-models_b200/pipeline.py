"""Pipeline-parallel LoRA fine-tuning of an OPT-style decoder over the GPUs of one box.

What the reference intends at finetune_lora_opt_pp.py:114-224 (it never ran: every committed log
ends in an OOM kill before step 0 and `batch_iter` is undefined): the N decoder
layers are split `N // S (+1)` per stage (finetune_lora_opt_pp.py:156), embeddings live on the
first stage (:121-130), final LayerNorm + lm_head + shifted cross-entropy on the last (:138-153),
micro-batches flow through the stages, AdamW + global-norm clip 1.0 on the adapter parameters
(:193-197).  DeepSpeed's PipelineModule/engine is absent here; this module restates that intent with
NCCL send/recv between adjacent ranks (gloo in the CPU tests) and a 1F1B schedule:

  * stage s runs min(S-1-s, m) warm-up forwards, then alternates one forward / one backward, then
    drains the remaining backwards; bubble fraction (S-1)/(m+S-1);
  * what crosses a stage boundary is the hidden state [mb, seq, hidden] forward and its gradient
    backward -- nothing else (the reference threads mask and labels through every stage,
    :130,136; only the last stage needs the labels, and the all-ones mask is rebuilt locally);
  * every adapter lives on exactly one stage, so there is no gradient all-reduce; the clip norm is
    the only cross-stage reduction (one all-reduce of a scalar per optimiser step).

Correctness criterion: a pipeline is a re-ordering, so loss and adapter gradients must equal those
of the unsplit model accumulated over the same micro-batches (tests/test_pipeline_cpu.py).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


def split_layers(n_layers: int, n_stages: int) -> List[int]:
    """finetune_lora_opt_pp.py:156."""
    return [n_layers // n_stages + (1 if i < n_layers % n_stages else 0) for i in range(n_stages)]


class OPTStage(nn.Module):
    """Layers [lo, hi) of an OPT decoder, plus the embeddings on the first stage and the
    final LayerNorm / lm_head / loss on the last one."""

    def __init__(self, causal_lm: nn.Module, stage: int, n_stages: int, loss_fn=None, ignore_index: int = -100):
        super().__init__()
        # ignore_index: label value that does not count in the loss.  -100 is what HF's collator writes over padding
        # (finetune_lora_distilgpt2.py:114); the reference's pipeline head instead feeds token ids as labels and
        # ignores tok.pad_token_id (finetune_lora_opt_pp.py:143-152): pass that id here to follow it.
        self.ignore_index = int(ignore_index)
        # loss_fn(logits, labels): default torch fp32 shifted cross-entropy; on the GPU pass
        # frozen.fused_causal_lm_loss (and swap lm_head / frozen neighbours with frozen.accelerate_* first).
        self.loss_fn = loss_fn
        decoder = causal_lm.model.decoder  # peft_model.base_model.model.model.decoder, :116-118
        counts = split_layers(len(decoder.layers), n_stages)
        lo = sum(counts[:stage])
        self.stage, self.n_stages = stage, n_stages
        self.is_first, self.is_last = stage == 0, stage == n_stages - 1
        self.layers = nn.ModuleList(list(decoder.layers[lo:lo + counts[stage]]))
        self.config = causal_lm.config
        if self.is_first:
            self.embed_tokens = decoder.embed_tokens
            self.embed_positions = decoder.embed_positions
            self.project_in = decoder.project_in
        if self.is_last:
            self.final_layer_norm = decoder.final_layer_norm
            self.project_out = decoder.project_out
            self.lm_head = causal_lm.lm_head
        self.hidden_size = causal_lm.config.hidden_size

    def _causal_mask(self, hidden: torch.Tensor, attention_mask: torch.Tensor):
        from transformers.masking_utils import create_causal_mask

        return create_causal_mask(config=self.config, inputs_embeds=hidden, attention_mask=attention_mask,
                                  past_key_values=None)

    def forward(self, x: torch.Tensor, attention_mask: torch.Tensor, labels: Optional[torch.Tensor] = None):
        """x: input_ids [mb, seq] on the first stage, hidden [mb, seq, hidden] elsewhere.
        Returns hidden, or the scalar loss on the last stage when labels are given."""
        position_ids = (torch.cumsum(attention_mask, dim=1) * attention_mask - 1).long()
        if self.is_first:
            h = self.embed_tokens(x)
            pos = self.embed_positions(attention_mask, 0, position_ids=position_ids)
            if self.project_in is not None:
                h = self.project_in(h)
            h = h + pos.to(h.dtype)
        else:
            h = x
        mask = self._causal_mask(h, attention_mask)
        for layer in self.layers:
            h = layer(h, attention_mask=mask, position_ids=position_ids, past_key_values=None, use_cache=False)
            if isinstance(h, tuple):
                h = h[0]
        if not self.is_last:
            return h
        if self.final_layer_norm is not None:
            h = self.final_layer_norm(h)
        if self.project_out is not None:
            h = self.project_out(h)
        logits = self.lm_head(h)
        if labels is None:
            return logits
        # teacher-forcing next-token loss, finetune_lora_opt_pp.py:147-152
        if self.loss_fn is not None:
            return self.loss_fn(logits, labels, ignore_index=self.ignore_index)
        return F.cross_entropy(logits[:, :-1].contiguous().float().view(-1, logits.size(-1)),
                               labels[:, 1:].contiguous().view(-1), ignore_index=self.ignore_index)


class PipelineEngine:
    """`train_batch(micro_batches)` = one optimiser step over m micro-batches, 1F1B."""

    def __init__(self, stage_module: OPTStage, group=None, lr: float = 5e-5, max_grad_norm: float = 1.0,
                 act_dtype: Optional[torch.dtype] = None, device: Optional[torch.device] = None, overlap: bool = True):
        self.m = stage_module
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        assert self.world == stage_module.n_stages and self.rank == stage_module.stage
        self.params = [p for p in stage_module.parameters() if p.requires_grad]
        # AdamW betas (0.9, 0.999), eps 1e-8, wd 0, constant lr: finetune_lora_opt_pp.py:193-196, 207
        self.opt = torch.optim.AdamW(self.params, lr=lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0) \
            if self.params else None
        self.max_grad_norm = max_grad_norm
        p0 = next(stage_module.parameters())
        self.device = device or p0.device
        self.act_dtype = act_dtype or p0.dtype
        self.prev = self.rank - 1 if self.rank > 0 else None
        self.next = self.rank + 1 if self.rank < self.world - 1 else None
        # overlap: activations and gradients travel on two communicators of their own (one per direction), each
        # send is enqueued without waiting and only a receive makes the compute stream wait for data.  With one
        # communicator the two directions between a pair of neighbours share an ordered queue, which is why the
        # serial mode (overlap=False) must group "send forward + receive backward" into one blocking exchange.
        self.overlap = bool(overlap) and self.world > 1
        self.g_fwd = self.g_bwd = None
        if self.overlap:
            ranks = dist.get_process_group_ranks(group) if group is not None else list(range(self.world))
            self.g_fwd = dist.new_group(ranks=ranks)  # collective: every stage constructs its engine
            self.g_bwd = dist.new_group(ranks=ranks)
        self._inflight = []  # (work, tensor) of sends that have not been waited for
        self.p2p_bytes = 0   # bytes this stage sent + received in the last train_batch (for the bench)

    # ---- point-to-point -----------------------------------------------------------------
    # Each transfer is a one-element batch: under NCCL a batched P2P op runs on the pair's own communicator,
    # whereas an unbatched isend/irecv on an eagerly initialised group is queued behind every other op of the group.
    def _send(self, t, peer, g):
        t = t.contiguous()
        for w in dist.batch_isend_irecv([dist.P2POp(dist.isend, t, peer, g)]):
            self._inflight.append((w, t))
        self.p2p_bytes += t.numel() * t.element_size()

    def _recv(self, shape, peer, g):
        buf = torch.empty(shape, dtype=self.act_dtype, device=self.device)
        for w in dist.batch_isend_irecv([dist.P2POp(dist.irecv, buf, peer, g)]):
            w.wait()  # NCCL: the compute stream waits, the host does not
        self.p2p_bytes += buf.numel() * buf.element_size()
        return buf

    def _p2p(self, send_fwd=None, send_bwd=None, recv_fwd_shape=None, recv_bwd_shape=None):
        """The exchanges with both neighbours that are due at one point of the schedule."""
        if self.overlap:
            if send_bwd is not None:
                self._send(send_bwd, self.prev, self.g_bwd)
            if send_fwd is not None:
                self._send(send_fwd, self.next, self.g_fwd)
            out_f = self._recv(recv_fwd_shape, self.prev, self.g_fwd) if recv_fwd_shape is not None else None
            out_b = self._recv(recv_bwd_shape, self.next, self.g_bwd) if recv_bwd_shape is not None else None
            return out_f, out_b
        # serial mode: one grouped exchange (ncclGroupStart/End under NCCL), waited for as a whole
        ops, out_f, out_b = [], None, None
        if send_bwd is not None:
            ops.append(dist.P2POp(dist.isend, send_bwd.contiguous(), self.prev, self.group))
        if recv_fwd_shape is not None:
            out_f = torch.empty(recv_fwd_shape, dtype=self.act_dtype, device=self.device)
            ops.append(dist.P2POp(dist.irecv, out_f, self.prev, self.group))
        if send_fwd is not None:
            ops.append(dist.P2POp(dist.isend, send_fwd.contiguous(), self.next, self.group))
        if recv_bwd_shape is not None:
            out_b = torch.empty(recv_bwd_shape, dtype=self.act_dtype, device=self.device)
            ops.append(dist.P2POp(dist.irecv, out_b, self.next, self.group))
        if ops:
            for op in ops:
                self.p2p_bytes += op.tensor.numel() * op.tensor.element_size()
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        return out_f, out_b

    # ---- one optimiser step ---------------------------------------------------------------
    def train_batch(self, micro_batches: List[Dict[str, torch.Tensor]]) -> Optional[torch.Tensor]:
        """Every rank passes the same list (ids/mask/labels); only the first stage reads the ids
        and only the last reads the labels.  Returns the mean loss on the last stage.

        Exchanges with the two neighbours that are due at the same point of the schedule are
        issued as ONE group (send-forward + receive-backward after a forward, send-backward +
        receive-forward after a backward): neither stage can then sit in a send that its
        neighbour will only match after its own send completes."""
        m = len(micro_batches)
        S, s = self.world, self.rank
        self.p2p_bytes = 0
        mb, seq = micro_batches[0]["input_ids"].shape
        hshape = (mb, seq, self.m.hidden_size)
        warm = min(S - 1 - s, m)
        steady = m - warm
        first, last = self.m.is_first, self.m.is_last
        inputs, outputs, losses = [], [], []
        self.m.train()
        n_fwd = [0]

        def fwd(x):
            b = micro_batches[n_fwd[0]]
            n_fwd[0] += 1
            mask = b["attention_mask"].to(self.device)
            if first:
                x = b["input_ids"].to(self.device)
            else:
                x.requires_grad_(True)
            out = self.m(x, mask, b["labels"].to(self.device) if last else None)
            if last:
                out = out / m
                losses.append(out.detach())
            inputs.append(x)
            outputs.append(out)
            return out

        def bwd(g):
            x, out = inputs.pop(0), outputs.pop(0)
            if last:
                out.backward()
            else:
                torch.autograd.backward(out, grad_tensors=g)
            return None if first else x.grad

        recv_f = None if first else hshape
        recv_b = None if last else hshape
        # warm-up: forwards only
        for _ in range(warm):
            x, _ = self._p2p(recv_fwd_shape=recv_f)
            out = fwd(x)
            self._p2p(send_fwd=None if last else out.detach())
        # steady state: one forward, one backward per iteration
        x = None
        if steady > 0:
            x, _ = self._p2p(recv_fwd_shape=recv_f)
        for i in range(steady):
            out = fwd(x)
            _, g = self._p2p(send_fwd=None if last else out.detach(), recv_bwd_shape=recv_b)
            gx = bwd(g)
            if i == steady - 1:
                self._p2p(send_bwd=gx)
            else:
                x, _ = self._p2p(send_bwd=gx, recv_fwd_shape=recv_f)
        # cool-down: drain the backwards of the warm-up forwards
        for _ in range(warm):
            _, g = self._p2p(recv_bwd_shape=recv_b)
            gx = bwd(g)
            self._p2p(send_bwd=gx)

        for w, _ in self._inflight:  # every send has left (NCCL: stream-level wait) before its buffer can be reused
            w.wait()
        self._inflight.clear()
        # global gradient-norm clip across stages (each adapter lives on exactly one stage)
        sq = torch.zeros((), dtype=torch.float32, device=self.device)
        for p in self.params:
            if p.grad is not None:
                sq += p.grad.detach().float().pow(2).sum()
        if self.world > 1:
            dist.all_reduce(sq, group=self.group)
        total = sq.sqrt()
        if self.max_grad_norm is not None and self.max_grad_norm > 0:
            coef = torch.clamp(self.max_grad_norm / (total + 1e-6), max=1.0)
            for p in self.params:
                if p.grad is not None:
                    p.grad.mul_(coef.to(p.grad.dtype))
        self.last_grad_norm = total
        if self.opt is not None:
            self.opt.step()
        return torch.stack(losses).sum() if self.m.is_last else None

    def zero_grad(self):
        if self.opt is not None:
            self.opt.zero_grad(set_to_none=True)


def save_adapters(stage: OPTStage, save_dir: str, peft_config=None, meta: Optional[dict] = None,
                  adapter_name: str = "default", group=None) -> None:
    """Adapter save for a pipeline run (finetune_lora_opt_pp.py:226-243: rank 0 writes the adapters and
    meta.json with the split).  Every adapter lives on exactly one stage, so each rank hands its own
    tensors -- renamed from the stage's local `layers.<i>` to the peft key of the unsplit model -- to
    rank 0, which writes ONE peft-format directory: adapter_model.safetensors, adapter_config.json (when
    the LoraConfig is given) and meta.json."""
    import json
    import os

    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    n_layers = stage.config.num_hidden_layers
    counts = split_layers(n_layers, stage.n_stages)
    lo = sum(counts[:stage.stage])
    mine = {}
    for name, prm in stage.named_parameters():
        if "lora_" not in name or f".{adapter_name}." not in name:
            continue
        local, rest = name.split(".", 2)[1:]
        key = f"base_model.model.model.decoder.layers.{lo + int(local)}.{rest}".replace(f".{adapter_name}.", ".")
        mine[key] = prm.detach().to("cpu", torch.float32).contiguous()
    shards = [None] * world
    if world > 1:
        dist.gather_object(mine, shards if rank == 0 else None, dst=0, group=group)
    else:
        shards = [mine]
    if rank == 0:
        merged = {}
        for s in shards:
            dup = set(merged) & set(s)
            if dup:
                raise RuntimeError(f"adapter tensors owned by two stages: {sorted(dup)[:3]}")
            merged.update(s)
        os.makedirs(save_dir, exist_ok=True)
        from safetensors.torch import save_file

        save_file(merged, os.path.join(save_dir, "adapter_model.safetensors"), metadata={"format": "pt"})
        if peft_config is not None:
            cfg = peft_config.to_dict()
            cfg["inference_mode"] = True
            with open(os.path.join(save_dir, "adapter_config.json"), "w") as f:
                json.dump(cfg, f, indent=2, sort_keys=True)
        m = {"split": counts, "stages": stage.n_stages}
        m.update(meta or {})
        with open(os.path.join(save_dir, "meta.json"), "w") as f:
            json.dump(m, f, indent=2)
    if world > 1:
        dist.barrier(group=group)
