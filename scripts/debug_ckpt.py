import os, sys, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
import __graft_entry__ as g
pkg = g.load_package()
import test_gpu_model as T
log = []
orig = pkg.lora._philox_seed_offset
def spy(device, numel):
    s, o = orig(device, numel)
    log.append((s, o, torch.is_grad_enabled()))
    return s, o
pkg.lora._philox_seed_offset = spy
base = T._gpt2()
cfg = pkg.LoraConfig(r=8, lora_alpha=16, target_modules=["c_attn", "c_proj"], lora_dropout=0.05, task_type=pkg.TaskType.CAUSAL_LM)
ids = torch.randint(0, 512, (2, 64), generator=torch.Generator().manual_seed(5)).cuda()
grads = []
for ckpt in (False, True):
    log.clear()
    m = copy.deepcopy(base).to(torch.bfloat16)
    if ckpt: m.gradient_checkpointing_enable()
    m.config.use_cache = False
    pm = pkg.get_peft_model(m, cfg).cuda()
    with torch.no_grad():
        for n, p in pm.named_parameters():
            if "lora_B" in n: p.copy_(torch.randn(p.shape, generator=torch.Generator().manual_seed(11)).to(p.device) * 0.02)
    pm.train()
    torch.manual_seed(123)
    loss = pkg.train.causal_lm_loss(pm(input_ids=ids).logits, ids)
    nf = len(log)
    loss.backward()
    print("ckpt", ckpt, "loss", float(loss), "fwd calls", nf, "total calls", len(log))
    print("  fwd   :", [(o, ge) for _, o, ge in log[:nf]])
    print("  in bwd:", [(o, ge) for _, o, ge in log[nf:]])
    grads.append({n: p.grad.clone() for n, p in pm.named_parameters() if p.requires_grad})
for n in grads[0]:
    print(round(T.relerr(grads[1][n], grads[0][n]), 5), n)
