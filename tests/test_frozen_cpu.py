"""Host-side logic of the callers either side of the path (no GPU): module surgery, weight tying,
loud failure without CUDA (there is no CPU fallback anywhere in the product)."""
import pytest
import torch
import torch.nn as nn


def _gpt2(pkg):
    from transformers import GPT2Config, GPT2LMHeadModel

    m = GPT2LMHeadModel(GPT2Config(n_layer=2, n_embd=64, n_head=2, vocab_size=101, n_positions=32))
    return pkg.get_peft_model(m, pkg.LoraConfig(r=8, target_modules=["c_attn", "c_proj"], task_type=pkg.TaskType.CAUSAL_LM))


def _opt(pkg, targets):
    from transformers import OPTConfig, OPTForCausalLM

    m = OPTForCausalLM(OPTConfig(num_hidden_layers=2, hidden_size=64, ffn_dim=128, num_attention_heads=2, vocab_size=104,
                                 word_embed_proj_dim=64, max_position_embeddings=32))
    return pkg.get_peft_model(m, pkg.LoraConfig(r=8, target_modules=targets, task_type=pkg.TaskType.CAUSAL_LM))


def test_gpt2_neighbours_surgery(pkg):
    pm = _gpt2(pkg)
    assert pkg.frozen.accelerate_frozen_neighbours(pm) == 2
    for blk in pm.base_model.model.transformer.h:
        assert isinstance(blk.mlp.c_fc, pkg.frozen.FrozenLinear) and blk.mlp.c_fc.act == "gelu_new"
        assert blk.mlp.c_fc.w_is_KN == 1 and (blk.mlp.c_fc.in_features, blk.mlp.c_fc.out_features) == (64, 256)
        assert isinstance(blk.mlp.act, nn.Identity)
        assert isinstance(blk.mlp.c_proj, pkg.lora.Linear) and isinstance(blk.attn.c_attn, pkg.lora.Linear)
    assert pkg.frozen.accelerate_frozen_neighbours(pm) == 0  # idempotent
    # trainable set unchanged: adapters only
    assert all(("lora_" in n) == p.requires_grad for n, p in pm.named_parameters())


def test_opt_neighbours_surgery_respects_targets(pkg):
    pm = _opt(pkg, ["q_proj", "v_proj"])
    assert pkg.frozen.accelerate_frozen_neighbours(pm, plain=True) == 2 * 4  # fc1+relu, k_proj, out_proj, fc2 per layer
    lyr = pm.base_model.model.model.decoder.layers[0]
    assert lyr.fc1.act == "relu" and isinstance(lyr.activation_fn, nn.Identity)
    assert isinstance(lyr.self_attn.q_proj, pkg.lora.Linear) and isinstance(lyr.self_attn.k_proj, pkg.frozen.FrozenLinear)
    pm6 = _opt(pkg, ["q_proj", "k_proj", "v_proj", "out_proj", "fc1", "fc2"])
    assert pkg.frozen.accelerate_frozen_neighbours(pm6, plain=True) == 0  # every neighbour is an adapter: nothing to do
    assert isinstance(pm6.base_model.model.model.decoder.layers[0].activation_fn, nn.ReLU)


def test_lm_head_swap_keeps_weight_tying(pkg):
    pm = _gpt2(pkg)
    pkg.frozen.accelerate_lm_head(pm)
    core = pm.base_model.model
    assert isinstance(core.lm_head, pkg.frozen.FrozenLinear)
    assert core.lm_head.weight is core.transformer.wte.weight
    with pytest.raises(ValueError):
        pkg.frozen.accelerate_lm_head(pm)  # no plain nn.Linear head left


def test_frozen_linear_refuses_trainable_layers_and_unknown_types(pkg):
    with pytest.raises(ValueError):
        pkg.frozen.FrozenLinear(nn.Linear(8, 8))
    with pytest.raises(ValueError):
        pkg.frozen.FrozenLinear(nn.Conv1d(8, 8, 1).requires_grad_(False))
    with pytest.raises(ValueError):
        pkg.frozen.FrozenLinear(nn.Linear(8, 8).requires_grad_(False), act="tanh")


def test_no_cpu_fallback(pkg):
    lin = nn.Linear(16, 16).requires_grad_(False)
    with pytest.raises(RuntimeError):
        pkg.frozen.FrozenLinear(lin)(torch.zeros(2, 16))
    with pytest.raises(RuntimeError):
        pkg.frozen.FrozenLinear(lin, act="relu")(torch.zeros(2, 16))
    with pytest.raises(RuntimeError):
        pkg.frozen.fused_causal_lm_loss(torch.zeros(1, 4, 8, dtype=torch.bfloat16), torch.zeros(1, 4, dtype=torch.long))
    with pytest.raises(RuntimeError):
        pkg.comm.PeerBucketExchange(1024)  # no process group / no CUDA
    with pytest.raises((RuntimeError, ValueError)):
        pkg.train.wrap_ddp(nn.Linear(4, 4), None, exchange="p2p")
