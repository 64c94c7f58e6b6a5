"""The adapter optimiser step on the device (lora_adamw_clip_step: global-norm clip + AdamW in two launches)
against the pair it replaces, torch.nn.utils.clip_grad_norm_ + torch.optim.AdamW -- what HF Trainer runs for the
reference (finetune_lora_distilgpt2.py:130-145) -- on the same parameters and gradients, fp32, several steps, with
tensors of ragged sizes (chunk tails), a parameter without gradient, weight decay on and off, clipping active and not."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("wd,max_norm,gscale", [(0.0, 1.0, 1.0), (0.01, 1.0, 1e-3), (0.0, 0.0, 1.0)])
def test_fused_adamw_matches_torch(pkg, wd, max_norm, gscale):
    g = torch.Generator().manual_seed(0)
    shapes = [(8, 768), (2304, 8), (8, 3072), (5000,), (1,), (4097,), (16, 2560), (3, 5)]
    ref_p = [torch.nn.Parameter(torch.randn(s, generator=g).cuda()) for s in shapes]
    our_p = [torch.nn.Parameter(p.detach().clone()) for p in ref_p]
    ref_opt = torch.optim.AdamW(ref_p, lr=1e-2, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd)
    ours = pkg.train.FusedAdamW(our_p, lr=1e-2, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd, max_grad_norm=max_norm)
    for step in range(5):
        skip = 3 if step == 2 else -1  # one step in which a parameter has no gradient
        for i, (a, b) in enumerate(zip(ref_p, our_p)):
            if i == skip:
                a.grad = b.grad = None
                continue
            gr = (torch.randn(a.shape, generator=g) * gscale).cuda()
            a.grad, b.grad = gr.clone(), gr.clone()
        want_norm = torch.nn.utils.clip_grad_norm_(ref_p, max_norm) if max_norm > 0 else \
            torch.norm(torch.stack([p.grad.norm() for p in ref_p if p.grad is not None]))
        ref_opt.step()
        got_norm = ours.step(lr=1e-2)
        torch.cuda.synchronize()
        assert abs(float(got_norm) - float(want_norm)) <= 1e-5 * float(want_norm)
        for i, (a, b) in enumerate(zip(ref_p, our_p)):
            assert torch.allclose(b.detach(), a.detach(), rtol=2e-5, atol=1e-7), (step, i)
            if a.grad is not None:  # clip_grad_norm_ scales the gradients in place; so does the kernel
                assert torch.allclose(b.grad, a.grad, rtol=1e-5, atol=1e-9), (step, i)


def test_fused_adamw_is_bit_reproducible(pkg):
    """Same inputs twice -> identical bits (the norm is folded in a fixed order, no floating-point atomics)."""
    outs = []
    for _ in range(2):
        g = torch.Generator().manual_seed(1)
        ps = [torch.nn.Parameter(torch.randn(n, generator=g).cuda()) for n in (100000, 33, 4096 * 3)]
        opt = pkg.train.FusedAdamW(ps, lr=1e-3)
        for _ in range(3):
            for p in ps:
                p.grad = torch.randn(p.shape, generator=g).cuda()
            opt.step()
        outs.append([p.detach().clone() for p in ps])
    assert all(torch.equal(a, b) for a, b in zip(*outs))


def test_fused_adamw_refuses_cpu_parameters(pkg):
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.train.FusedAdamW([torch.nn.Parameter(torch.zeros(4))])
