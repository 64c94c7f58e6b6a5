"""OPT-2.7B-architecture LoRA pipeline-parallel step over the GPUs of one box (BASELINE.json configs[4]):
fp16, 32 layers split over WORLD_SIZE stages, NCCL send/recv, 1F1B over --micro micro-batches.
    torchrun --nproc-per-node S scripts/pp_bench.py [--layers 32 --micro 8 --seq 2048 --steps 5]"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import __graft_entry__ as g

ap = argparse.ArgumentParser()
ap.add_argument("--layers", type=int, default=32)
ap.add_argument("--micro", type=int, default=8)
ap.add_argument("--mb", type=int, default=1)
ap.add_argument("--seq", type=int, default=2048)
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--warmup", type=int, default=2)
ap.add_argument("--stock-head", action="store_true", help="torch lm_head + fp32 cross-entropy on the last stage")
a = ap.parse_args()
g.build()
pkg = g.load_package()
rank, world, local = pkg.train.init_distributed("nccl")
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
from transformers import OPTConfig, OPTForCausalLM
cfg = OPTConfig(hidden_size=2560, ffn_dim=10240, num_hidden_layers=a.layers, num_attention_heads=32,
                word_embed_proj_dim=2560, vocab_size=50272, max_position_embeddings=2048, do_layer_norm_before=True,
                activation_function="relu", dropout=0.1)
torch.manual_seed(0)
# Build only this stage's share on the GPU: construct on meta, materialise the stage's modules.
with torch.device("meta"):
    full = OPTForCausalLM(cfg).to(torch.float16)
full.config.use_cache = False
pm = pkg.get_peft_model(full, pkg.LoraConfig(r=8, lora_alpha=16, lora_dropout=0.05, bias="none",
                                             target_modules=["q_proj", "k_proj", "v_proj", "out_proj", "fc1", "fc2"],
                                             task_type=pkg.TaskType.CAUSAL_LM))
loss_fn = None
if not a.stock_head:  # lm_head on the tcgen05 GEMM, fused cross-entropy (frozen.py); all six projections are adapters here
    pkg.frozen.accelerate_lm_head(pm)
    loss_fn = pkg.frozen.fused_causal_lm_loss
stage = pkg.pipeline.OPTStage(pm.base_model.model, rank, world, loss_fn=loss_fn)
stage.to_empty(device=dev)
gen = torch.Generator(device=dev).manual_seed(1 + rank)
with torch.no_grad():
    for n, p in stage.named_parameters():
        if "lora_B" in n:
            p.zero_()
        elif p.dim() > 1:
            p.normal_(0.0, 0.02, generator=gen)
        elif "layer_norm" in n and n.endswith("weight"):
            p.fill_(1.0)
        else:
            p.zero_()
for n, p in stage.named_parameters():
    p.requires_grad_("lora_" in n)
eng = pkg.pipeline.PipelineEngine(stage, lr=5e-5, max_grad_norm=1.0, act_dtype=torch.float16, device=dev)
batches = pkg.train.synthetic_batches(a.micro, a.mb, a.seq, cfg.vocab_size, 1, seed=1234, device=dev)

def step():
    eng.zero_grad()
    return eng.train_batch(batches)

for _ in range(a.warmup):
    step()
dist.barrier(); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.steps):
    loss = step()
e1.record()
dist.barrier(); torch.cuda.synchronize()
t = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
l = torch.tensor([float(loss) if loss is not None else 0.0], device=dev)
dist.all_reduce(l)
if rank == 0:
    tokens = a.micro * a.mb * a.seq
    print(json.dumps({"metric": "lora_finetune_tokens_per_sec", "value": tokens / (float(t) * 1e-3), "unit": "tokens/s",
                      "n_gpus": world, "ms_per_step": float(t), "dtype": "fp16", "loss": float(l),
                      "config": {"workload": f"OPT-2.7B-arch LoRA (6 targets, r=8) pipeline-parallel, {a.layers} layers over "
                                             f"{world} stages, 1F1B, {a.micro} micro-batches of {a.mb} x {a.seq}",
                                 "split": pkg.pipeline.split_layers(a.layers, world), "parallelism": f"pp{world}"}}))
dist.destroy_process_group()
