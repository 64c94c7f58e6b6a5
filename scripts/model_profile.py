"""GPU-time breakdown of one whole-model LoRA step (torch profiler): which kernels bound tokens/s."""
import os, sys, argparse
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as g
import bench

ap = argparse.ArgumentParser(); ap.add_argument("--workload", default="distilgpt2"); ap.add_argument("--no-ckpt", action="store_true"); ap.add_argument("--stock-head", action="store_true")
args = ap.parse_args()
g.build(); pkg = g.load_package()
a = argparse.Namespace(workload=args.workload, layers=0, batch=0, seq=0, dtype="bf16", no_ckpt=args.no_ckpt)
cls, cfg, targets, B, S, pad_id = bench.model_spec(a)
dev = torch.device("cuda:0")
torch.manual_seed(0)
with torch.device(dev):
    model = cls(cfg).to(torch.bfloat16)
model.config.use_cache = False
if not args.no_ckpt: model.gradient_checkpointing_enable()
pm = pkg.get_peft_model(model, pkg.LoraConfig(r=8, lora_alpha=16, target_modules=targets, lora_dropout=0.05, task_type=pkg.TaskType.CAUSAL_LM))
loss_fn = None
if not args.stock_head:
    pkg.frozen.accelerate_lm_head(pm); pkg.frozen.accelerate_frozen_neighbours(pm); loss_fn = pkg.frozen.fused_causal_lm_loss
tr = pkg.train.Trainer(pm, pkg.train.TrainConfig(lr=5e-5, accum=1, total_steps=10**6), loss_fn=loss_fn); tr._ensure_sched(10**6); pm.train()
batches = pkg.train.synthetic_batches(2, B, S, cfg.vocab_size, pad_id, seed=1, device=dev)
def step(i):
    tr.micro_step(batches[i % 2], sync=True); tr.optimizer_step()
for i in range(3): step(i)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for i in range(3): step(i)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=40, max_name_column_width=70))
