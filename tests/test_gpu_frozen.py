"""SURVEY.md 8f rows N2 / N3 on the GPU: the un-adapted lm_head through the tcgen05 GEMM (ragged
vocab, padded row stride) and the fused shifted cross-entropy, against the CPU oracle
(torch fp32 F.linear / oracle.lora_ref.causal_lm_loss) on the same 16-bit-rounded inputs."""
import copy

import pytest
import torch

from oracle import lora_ref

pytestmark = pytest.mark.gpu
TOL = 2e-2


def relerr(a, ref):
    a, ref = a.detach().float().cpu(), ref.detach().float().cpu()
    return float((a - ref).abs().max() / ref.abs().max().clamp_min(1e-20))


@pytest.mark.parametrize("M,K,N,dtype", [(300, 256, 1003, torch.bfloat16), (1024, 768, 50257, torch.bfloat16),
                                         (130, 768, 512, torch.float16)])
def test_frozen_linear_ragged_vocab(pkg, M, K, N, dtype):
    g = torch.Generator().manual_seed(0)
    x = torch.randn(M, K, generator=g).to(dtype)
    W = (torch.randn(N, K, generator=g) * 0.02).to(dtype)
    dy = torch.randn(M, N, generator=g).to(dtype)
    lin = torch.nn.Linear(K, N, bias=False).to(dtype)
    with torch.no_grad():
        lin.weight.copy_(W)
    lin.requires_grad_(False)
    fl = pkg.frozen.FrozenLinear(lin.cuda())
    xg = x.cuda().requires_grad_(True)
    y = fl(xg)
    assert y.shape == (M, N)
    y.backward(dy.cuda())
    ref_y = x.float() @ W.float().t()
    ref_dx = dy.float() @ W.float()
    assert relerr(y, ref_y) < TOL and relerr(xg.grad, ref_dx) < TOL
    assert lin.weight.grad is None


@pytest.mark.parametrize("B,S,V,dtype", [(2, 33, 1003, torch.bfloat16), (4, 128, 50257, torch.bfloat16),
                                         (3, 17, 512, torch.float16)])
def test_fused_causal_lm_loss_matches_oracle(pkg, B, S, V, dtype):
    g = torch.Generator().manual_seed(1)
    logits = (torch.randn(B, S, V, generator=g) * 2).to(dtype)
    labels = torch.randint(0, V, (B, S), generator=g)
    labels[0, 3:6] = -100  # padded positions, DataCollatorForLanguageModeling style
    ref_in = logits.float().requires_grad_(True)
    ref = lora_ref.causal_lm_loss(ref_in, labels)
    (ref * 0.5).backward()
    lg = logits.cuda().requires_grad_(True)
    loss = pkg.frozen.fused_causal_lm_loss(lg, labels.cuda())
    (loss * 0.5).backward()
    assert abs(float(loss) - float(ref)) / abs(float(ref)) < 2e-3
    assert lg.grad.shape == (B, S, V)
    assert relerr(lg.grad, ref_in.grad) < TOL
    assert float(lg.grad[:, -1].float().abs().max()) == 0.0  # the last position has no target
    assert float(lg.grad[0, 2:5].float().abs().max()) == 0.0  # rows whose shifted label is -100


@pytest.mark.parametrize("M,K,N,act,conv1d,dtype", [
    (300, 256, 1024, "gelu_new", True, torch.bfloat16),
    (4096, 768, 3072, "gelu_new", True, torch.bfloat16),
    (520, 512, 2048, "relu", False, torch.float16),
    (130, 256, 264, "relu", False, torch.bfloat16),
    (130, 256, 264, None, True, torch.bfloat16),
])
def test_frozen_linear_activation_epilogue(pkg, M, K, N, act, conv1d, dtype):
    """Frozen c_fc + gelu_new (GPT-2 Conv1D layout) / fc1 + relu against the fp32 torch formula."""
    from transformers.activations import NewGELUActivation
    from transformers.pytorch_utils import Conv1D

    g = torch.Generator().manual_seed(4)
    x = torch.randn(M, K, generator=g).to(dtype)
    W = (torch.randn(N, K, generator=g) * 0.05).to(dtype)
    b = (torch.randn(N, generator=g) * 0.1).to(dtype)
    dy = torch.randn(M, N, generator=g).to(dtype)
    if conv1d:
        base = Conv1D(N, K)
        with torch.no_grad():
            base.weight.copy_(W.t().float())
            base.bias.copy_(b.float())
    else:
        base = torch.nn.Linear(K, N)
        with torch.no_grad():
            base.weight.copy_(W.float())
            base.bias.copy_(b.float())
    base = base.to(dtype).requires_grad_(False).cuda()
    fl = pkg.frozen.FrozenLinear(base, act=act)
    xg = x.cuda().requires_grad_(True)
    y = fl(xg)
    y.backward(dy.cuda())
    xr = x.float().requires_grad_(True)
    z = xr @ W.float().t() + b.float()
    ref = {"gelu_new": NewGELUActivation(), "relu": torch.nn.ReLU(), None: torch.nn.Identity()}[act](z)
    ref.backward(dy.float())
    assert relerr(y, ref) < TOL and relerr(xg.grad, xr.grad) < TOL
    with torch.no_grad():  # inference path: no pre-activation buffer
        assert relerr(fl(x.cuda()), ref) < TOL


def test_opt_layer_with_frozen_neighbours_matches_stock_path(pkg):
    from transformers import OPTConfig, OPTForCausalLM

    torch.manual_seed(0)
    base = OPTForCausalLM(OPTConfig(num_hidden_layers=2, hidden_size=256, ffn_dim=1024, num_attention_heads=4,
                                    vocab_size=1000, word_embed_proj_dim=256, max_position_embeddings=64,
                                    dropout=0.0)).to(torch.bfloat16)
    cfg = pkg.LoraConfig(r=8, lora_alpha=16, target_modules=["q_proj", "v_proj"], task_type=pkg.TaskType.CAUSAL_LM)
    ids = torch.randint(2, 1000, (2, 48), generator=torch.Generator().manual_seed(2)).cuda()
    res = []
    for fast in (False, True):
        torch.manual_seed(5)
        pm = pkg.get_peft_model(copy.deepcopy(base), cfg).cuda()
        with torch.no_grad():
            for n, p in pm.named_parameters():
                if "lora_B" in n:
                    p.copy_(torch.randn(p.shape, generator=torch.Generator().manual_seed(3)).to(p.device) * 0.02)
        if fast:
            assert pkg.frozen.accelerate_frozen_neighbours(pm, plain=True) == 2 * 4
            pkg.frozen.accelerate_lm_head(pm)
        pm.train()
        logits = pm(input_ids=ids).logits
        loss = pkg.frozen.fused_causal_lm_loss(logits, ids) if fast else pkg.train.causal_lm_loss(logits, ids)
        loss.backward()
        res.append((float(loss), {n: p.grad.clone() for n, p in pm.named_parameters() if p.requires_grad}))
    assert abs(res[0][0] - res[1][0]) / res[0][0] < 5e-3
    for n in res[0][1]:
        a, b = res[1][1][n].float().flatten(), res[0][1][n].float().flatten()
        assert float(torch.nn.functional.cosine_similarity(a, b, dim=0)) > 0.98, n  # bf16 + relu: see test_gpu_model


@pytest.mark.parametrize("B,S,K,V,chunk,dtype", [(2, 33, 64, 1003, 16, torch.bfloat16),  # ragged vocab, ragged last chunk
                                                 (4, 128, 768, 50257, 256, torch.bfloat16),  # distilgpt2 head
                                                 (3, 17, 128, 512, 0, torch.float16)])  # default chunk (one chunk)
def test_chunked_head_loss_matches_oracle(pkg, B, S, K, V, chunk, dtype):
    """lm_head + shifted cross-entropy in token chunks (no [M, V] logits): loss and d(loss)/d(hidden) against the
    fp32 CPU formula F.linear -> lora_ref.causal_lm_loss on the same 16-bit-rounded operands
    (finetune_lora_distilgpt2.py:114, finetune_lora_opt_pp.py:143-153)."""
    g = torch.Generator().manual_seed(4)
    hidden = torch.randn(B, S, K, generator=g).to(dtype)
    W = (torch.randn(V, K, generator=g) * 0.05).to(dtype)
    labels = torch.randint(0, V, (B, S), generator=g)
    labels[0, 3:6] = -100
    h_ref = hidden.float().requires_grad_(True)
    ref = lora_ref.causal_lm_loss(torch.nn.functional.linear(h_ref, W.float()), labels)
    (ref * 0.5).backward()
    h = hidden.cuda().requires_grad_(True)
    loss = pkg.frozen.chunked_head_loss(h, W.cuda(), labels.cuda(), chunk_tokens=chunk)
    (loss * 0.5).backward()
    assert abs(float(loss) - float(ref)) / abs(float(ref)) < 2e-3
    assert h.grad.shape == (B, S, K)
    assert relerr(h.grad, h_ref.grad) < TOL
    assert float(h.grad[:, -1].float().abs().max()) == 0.0  # the last position has no target
    # and against the two-kernel path that does materialise the logits
    h2 = hidden.cuda().requires_grad_(True)
    lin = torch.nn.Linear(K, V, bias=False).to(dtype).cuda()
    with torch.no_grad():
        lin.weight.copy_(W.cuda())
    lin.weight.requires_grad_(False)
    loss2 = pkg.frozen.fused_causal_lm_loss(pkg.frozen.FrozenLinear(lin)(h2), labels.cuda())
    (loss2 * 0.5).backward()
    assert abs(float(loss) - float(loss2)) < 1e-5 * abs(float(loss2)) + 1e-6
    assert relerr(h.grad, h2.grad) < 1e-2


def test_model_with_deferred_head_matches_materialised_path(pkg):
    """fuse_lm_head_with_loss on a GPT-2-style model: same loss and adapter gradients as lm_head + fused CE."""
    from transformers import GPT2Config, GPT2LMHeadModel

    torch.manual_seed(0)
    base = GPT2LMHeadModel(GPT2Config(n_layer=2, n_embd=256, n_head=4, vocab_size=1003, n_positions=64,
                                      resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)).to(torch.bfloat16)
    cfg = pkg.LoraConfig(r=8, lora_alpha=16, target_modules=["c_attn", "c_proj"], task_type=pkg.TaskType.CAUSAL_LM)
    ids = torch.randint(0, 1003, (2, 48), generator=torch.Generator().manual_seed(2)).cuda()
    res = []
    for deferred in (False, True):
        torch.manual_seed(5)
        pm = pkg.get_peft_model(copy.deepcopy(base), cfg).cuda()
        with torch.no_grad():
            for n, p in pm.named_parameters():
                if "lora_B" in n:
                    p.copy_(torch.randn(p.shape, generator=torch.Generator().manual_seed(3)).to(p.device) * 0.02)
        if deferred:
            loss_fn = pkg.frozen.fuse_lm_head_with_loss(pm, chunk_tokens=32)
            assert isinstance(pm.base_model.model.lm_head, pkg.frozen.DeferredHead)
        else:
            pkg.frozen.accelerate_lm_head(pm)
            loss_fn = pkg.frozen.fused_causal_lm_loss
        pm.train()
        loss = loss_fn(pm(input_ids=ids).logits, ids)
        loss.backward()
        res.append((float(loss), {n: p.grad.clone() for n, p in pm.named_parameters() if p.requires_grad}))
        if deferred:  # eval mode: the deferred head is the plain frozen projection again
            pm.eval()
            with torch.no_grad():
                assert pm(input_ids=ids).logits.shape == (2, 48, 1003)
    assert abs(res[0][0] - res[1][0]) / res[0][0] < 1e-3
    for n in res[0][1]:
        assert relerr(res[1][1][n], res[0][1][n]) < 2e-2, n


def test_model_with_fast_head_and_loss_matches_stock_path(pkg):
    """distilgpt2-style model with a ragged vocab: swapping lm_head + loss changes nothing but speed."""
    from transformers import GPT2Config, GPT2LMHeadModel

    torch.manual_seed(0)
    base = GPT2LMHeadModel(GPT2Config(n_layer=2, n_embd=256, n_head=4, vocab_size=1003, n_positions=64,
                                      resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)).to(torch.bfloat16)
    cfg = pkg.LoraConfig(r=8, lora_alpha=16, target_modules=["c_attn", "c_proj"], task_type=pkg.TaskType.CAUSAL_LM)
    ids = torch.randint(0, 1003, (2, 48), generator=torch.Generator().manual_seed(2)).cuda()
    res = []
    for fast in (False, True):
        torch.manual_seed(5)
        pm = pkg.get_peft_model(copy.deepcopy(base), cfg).cuda()
        with torch.no_grad():
            for n, p in pm.named_parameters():
                if "lora_B" in n:
                    p.copy_(torch.randn(p.shape, generator=torch.Generator().manual_seed(3)).to(p.device) * 0.02)
        if fast:
            pkg.frozen.accelerate_lm_head(pm)
            assert isinstance(pm.base_model.model.lm_head, pkg.frozen.FrozenLinear)
            assert pkg.frozen.accelerate_frozen_neighbours(pm) == 2  # c_fc + gelu_new of both blocks
        pm.train()
        logits = pm(input_ids=ids).logits
        loss = pkg.frozen.fused_causal_lm_loss(logits, ids) if fast else pkg.train.causal_lm_loss(logits, ids)
        loss.backward()
        res.append((float(loss), {n: p.grad.clone() for n, p in pm.named_parameters() if p.requires_grad}))
    assert abs(res[0][0] - res[1][0]) / res[0][0] < 5e-3
    for n in res[0][1]:
        assert relerr(res[1][1][n], res[0][1][n]) < 5e-2, n


def test_disabled_or_merged_adapter_still_backpropagates_to_the_input(pkg):
    """peft semantics: with adapters disabled (or merged) the layer is the frozen base layer, and dX flows."""
    torch.manual_seed(0)
    base = torch.nn.Linear(256, 384).to(torch.bfloat16).cuda()
    layer = pkg.lora.Linear(base, "default", r=8, lora_alpha=16, lora_dropout=0.0)
    with torch.no_grad():
        layer.lora_B["default"].weight.normal_(0, 0.05)
    x = torch.randn(96, 256, device="cuda", dtype=torch.bfloat16)
    dy = torch.randn(96, 384, device="cuda", dtype=torch.bfloat16)
    W, b = base.weight.float().cpu(), base.bias.float().cpu()
    layer.enable_adapters(False)
    xg = x.clone().requires_grad_(True)
    y = layer(xg)
    y.backward(dy)
    assert relerr(y, x.float().cpu() @ W.t() + b) < TOL and relerr(xg.grad, dy.float().cpu() @ W) < TOL
    layer.enable_adapters(True)
    delta = (layer.lora_B["default"].weight.float() @ layer.lora_A["default"].weight.float()).cpu() * layer.scaling["default"]
    layer.merge()
    xg = x.clone().requires_grad_(True)
    y = layer(xg)
    y.backward(dy)
    Wm = W + delta
    assert relerr(y, x.float().cpu() @ Wm.t() + b) < TOL and relerr(xg.grad, dy.float().cpu() @ Wm) < TOL
    layer.unmerge()


@pytest.mark.parametrize("conv1d,dtype", [(False, torch.bfloat16), (True, torch.bfloat16), (True, torch.float16)])
def test_merge_kernel_both_layouts(pkg, conv1d, dtype):
    """W <- W + s B A in place (row N4): against the fp32 sum rounded once; unmerge returns W up to one rounding."""
    from transformers.pytorch_utils import Conv1D

    K, N, r = 264, 520, 8
    torch.manual_seed(3)
    base = (Conv1D(N, K) if conv1d else torch.nn.Linear(K, N)).to(dtype).cuda()
    with torch.no_grad():
        base.weight.normal_(0, 0.05)
    layer = pkg.lora.Linear(base, "default", r=r, lora_alpha=16, lora_dropout=0.0, fan_in_fan_out=conv1d)
    with torch.no_grad():
        layer.lora_B["default"].weight.normal_(0, 0.05)
    w0 = base.weight.detach().clone()
    delta = (layer.lora_B["default"].weight.float() @ layer.lora_A["default"].weight.float()) * layer.scaling["default"]
    if conv1d:
        delta = delta.t()
    layer.merge()
    want = (w0.float() + delta).to(dtype)
    ulp = 2.0 ** (-8 if dtype == torch.bfloat16 else -11)
    got = base.weight.detach()
    # fp32 product and sum, one rounding: equal to the reference except where fp32 summation order flips a rounding
    assert float((got != want).float().mean()) < 1e-3
    assert float((got.float() - want.float()).abs().max()) <= ulp * float(want.float().abs().max())
    layer.unmerge()
    assert float((base.weight.detach().float() - w0.float()).abs().max()) <= 2 * ulp * float(want.float().abs().max())


def test_committed_golden_vectors(pkg):
    """tests/golden/frozen_golden.npz (oracle/make_golden_frozen.py: transformers Conv1D + NewGELUActivation,
    nn.Linear + ReLU, shifted F.cross_entropy, all fp32 on the CPU) through the device kernels."""
    import os

    import numpy as np
    from transformers.pytorch_utils import Conv1D

    gold = np.load(os.path.join(os.path.dirname(__file__), "golden", "frozen_golden.npz"))
    T = lambda k: torch.from_numpy(gold[k])
    for name, act in (("gelu", "gelu_new"), ("relu", "relu")):
        M, K, N, kn = (int(v) for v in gold[f"{name}_meta"])
        base = Conv1D(N, K) if kn else torch.nn.Linear(K, N)
        with torch.no_grad():
            base.weight.copy_(T(f"{name}_W"))
            base.bias.copy_(T(f"{name}_b"))
        fl = pkg.frozen.FrozenLinear(base.to(torch.bfloat16).requires_grad_(False).cuda(), act=act)
        xg = T(f"{name}_x").to(torch.bfloat16).cuda().requires_grad_(True)
        y = fl(xg)
        y.backward(T(f"{name}_dy").to(torch.bfloat16).cuda())
        assert relerr(y, T(f"{name}_y")) < TOL and relerr(xg.grad, T(f"{name}_dx")) < TOL, name
    lg = T("ce_logits").to(torch.bfloat16).cuda().requires_grad_(True)
    loss = pkg.frozen.fused_causal_lm_loss(lg, T("ce_labels").cuda())
    loss.backward()
    assert abs(float(loss.detach()) - float(gold["ce_loss"])) / float(gold["ce_loss"]) < 2e-3
    assert relerr(lg.grad, T("ce_dlogits")) < TOL


def test_no_cpu_fallback(pkg):
    with pytest.raises(RuntimeError):
        pkg.frozen.fused_causal_lm_loss(torch.zeros(1, 4, 8), torch.zeros(1, 4, dtype=torch.long))


@pytest.mark.parametrize("M,K,N,conv1d,dtype,bias", [(1024, 2560, 2560, False, torch.bfloat16, True),
                                                     (300, 3072, 768, True, torch.bfloat16, True),
                                                     (64, 768, 520, False, torch.float16, False)])
def test_frozen_linear_residual_epilogue(pkg, M, K, N, conv1d, dtype, bias):
    """residual + linear(x) with the add in the GEMM epilogue (lora_frozen_linear_add_fwd) against torch in fp32:
    forward, dX and the residual's gradient; paired-CTA and single-CTA kernels, ragged N, Conv1D layout, fp16."""
    g = torch.Generator().manual_seed(3)
    x = torch.randn(M, K, generator=g).to(dtype).cuda().requires_grad_(True)
    W = (torch.randn((K, N) if conv1d else (N, K), generator=g) * 0.02).to(dtype).cuda()
    b = (torch.randn(N, generator=g) * 0.1).to(dtype).cuda() if bias else None
    R = torch.randn(M, N, generator=g).to(dtype).cuda().requires_grad_(True)
    y = pkg.frozen.frozen_linear_add(x, W, b, R, 1 if conv1d else 0)
    dy = torch.randn(M, N, generator=g).to(dtype).cuda()
    y.backward(dy)
    xf, Rf = x.detach().float().requires_grad_(True), R.detach().float().requires_grad_(True)
    Wf = W.float()
    ref = (xf @ Wf if conv1d else xf @ Wf.t()) + (b.float() if bias else 0.0) + Rf
    ref.backward(dy.float())

    def relerr(a, r):
        return float((a.float() - r).abs().max() / r.abs().max())

    assert relerr(y, ref.detach()) < 2e-2
    assert relerr(x.grad, xf.grad) < 2e-2
    assert torch.equal(R.grad, dy)


def test_frozen_linear_residual_in_place(pkg):
    """R may be Y itself: the C entry point accumulates a projection onto a buffer that already holds a term."""
    g = torch.Generator().manual_seed(4)
    M, K, N = 512, 768, 768
    x = torch.randn(M, K, generator=g).to(torch.bfloat16).cuda()
    W = (torch.randn(N, K, generator=g) * 0.02).to(torch.bfloat16).cuda()
    y = torch.randn(M, N, generator=g).to(torch.bfloat16).cuda()
    ref = x.float() @ W.float().t() + y.float()
    b = pkg.binding
    rc = b.lib().lora_frozen_linear_add_fwd(y.new_empty(0).data_ptr() or None, K, None, None, None, N, None, N, M, K, N,
                                            b.dtype_code(y.dtype), 0, None)
    assert rc != 0 and "null pointer" in b.last_error()
    import ctypes
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = b.lib().lora_frozen_linear_add_fwd(x.data_ptr(), K, W.data_ptr(), None, y.data_ptr(), N, y.data_ptr(), N, M, K, N,
                                            b.dtype_code(y.dtype), 0, st)
    assert rc == 0, b.last_error()
    torch.cuda.synchronize()
    assert float((y.float() - ref).abs().max() / ref.abs().max()) < 2e-2

