import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as g
pkg = g.load_package()
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import test_gpu_model as T
from oracle import lora_ref
for targets in (["q_proj"], ["k_proj"], ["v_proj"], ["out_proj"], ["fc1"], ["fc2"], ["q_proj","k_proj","v_proj","out_proj","fc1","fc2"]):
    ref, ours = T._pair(pkg, T._opt(), targets)
    ids = torch.randint(0, 512, (4, 64), generator=torch.Generator().manual_seed(3))
    ref.train(); ours.train()
    out_r = ref(input_ids=ids).logits; lora_ref.causal_lm_loss(out_r, ids).backward()
    out_o = ours(input_ids=ids.cuda()).logits; pkg.train.causal_lm_loss(out_o, ids.cuda()).backward()
    gr = {n: p.grad for n, p in ref.named_parameters() if p.requires_grad}
    go = {n: p.grad for n, p in ours.named_parameters() if p.requires_grad}
    errs = sorted(((T.relerr(go[n], gr[n]), n, float(gr[n].abs().max())) for n in gr), reverse=True)
    print(targets, "logits err", T.relerr(out_o, out_r))
    for e in errs[:4]: print("   ", e)
