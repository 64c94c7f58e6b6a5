"""BASELINE.json configs[0] / BASELINE.md section 2 "B-cfg1": the reference's own CPU-runnable case, timed.

distilgpt2 architecture (random init, GPT2Config(n_layer=6)) with LoRA r=8, alpha=16, dropout 0.05 on c_attn / c_proj
(18 modules, 405 504 trainable parameters), fp32, DDP over gloo with world_size 2, seq 128, batch 4 per rank, AdamW
lr 5e-5, clip 1.0 -- finetune_lora_distilgpt2.py:98-104, 130-145 -- on the oracle's layer (peft is not installable).
3 warm-up + 20 timed optimiser steps, then the 50-step loss curve.  Prints one JSON line on rank 0.

    OMP_NUM_THREADS=<cores/2> torchrun --nproc-per-node 2 --master-addr 127.0.0.1 scripts/b_cfg1_cpu.py
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import __graft_entry__ as g  # noqa: E402
from oracle import lora_ref  # noqa: E402


def main():
    from transformers import GPT2Config, GPT2LMHeadModel

    pkg = g.load_package()
    rank, world, _ = pkg.train.init_distributed("gloo")
    threads = int(os.environ.get("OMP_NUM_THREADS", "0")) or max(1, len(os.sched_getaffinity(0)) // max(world, 1))
    torch.set_num_threads(threads)
    B, S, warm, timed, curve = 4, 128, 3, 20, 50
    torch.manual_seed(0)  # finetune_lora_distilgpt2.py:30
    model = GPT2LMHeadModel(GPT2Config(n_layer=6, n_embd=768, n_head=12, n_positions=1024, vocab_size=50257))
    model.config.use_cache = False
    pm = lora_ref.ref_get_peft_model(model, lora_ref.RefLoraConfig(r=8, lora_alpha=16, lora_dropout=0.05,
                                                                  target_modules=["c_attn", "c_proj"]))
    n_train = sum(p.numel() for p in pm.parameters() if p.requires_grad)
    net = pkg.train.wrap_ddp(pm) if world > 1 else pm
    batches = pkg.train.synthetic_batches(warm + timed + curve, B, S, 50257, 50256, seed=1234 + rank)
    tr = pkg.train.Trainer(net, pkg.train.TrainConfig(lr=5e-5, accum=1, total_steps=warm + timed + curve))
    tr._ensure_sched(warm + timed + curve)
    net.train()

    def step(b):
        loss = tr.micro_step(b, sync=True)
        tr.optimizer_step()
        return float(loss)

    for b in batches[:warm]:
        step(b)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for b in batches[warm:warm + timed]:
        step(b)
    if world > 1:
        dist.barrier()
    wall = time.perf_counter() - t0
    losses = [step(b) for b in batches[warm + timed:]]
    if rank == 0:
        print(json.dumps({
            "impl": "reference", "metric": "lora_finetune_tokens_per_sec", "value": world * B * S * timed / wall,
            "unit": "tokens/s", "ms_per_step": wall / timed * 1e3, "steps": timed, "warmup": warm, "dtype": "f32",
            "data": "synthetic", "higher_is_better": True,
            "config": {"workload": "BASELINE.json configs[0]: distilgpt2 architecture LoRA (r=8, c_attn/c_proj) DDP on "
                                   f"CPU/gloo world_size={world}, seq {S}, batch {B} per rank, fp32, AdamW lr 5e-5, clip 1.0",
                       "trainable_params": n_train, "parallelism": f"dp{world} (gloo)"},
            "cpu_baseline": {"kind": "port", "cores": threads * world, "threads_per_rank": threads,
                             "sample": f"{timed} timed optimiser steps after {warm} warm-up"},
            "loss_curve_50_steps": [round(v, 5) for v in losses],
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
