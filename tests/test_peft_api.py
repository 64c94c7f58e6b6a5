"""Host-side mirror of the peft API the reference uses (construction runs on CPU; the layers'
forward needs the GPU and is covered in test_gpu_model.py)."""
import json
import os

import pytest
import torch
import torch.nn as nn


def _tiny_gpt2():
    from transformers import GPT2Config, GPT2LMHeadModel

    torch.manual_seed(0)
    return GPT2LMHeadModel(GPT2Config(n_layer=6, n_embd=768, n_head=12, vocab_size=512, n_positions=64))


def test_distilgpt2_injection_matches_reference_call_site(pkg):
    """finetune_lora_distilgpt2.py:98-104: r=8, alpha=16, ["c_attn","c_proj"], dropout .05 ->
    18 Conv1D modules wrapped, 405 504 trainable parameters, everything else frozen."""
    cfg = pkg.LoraConfig(r=8, lora_alpha=16, target_modules=["c_attn", "c_proj"], lora_dropout=0.05, bias="none",
                         task_type=pkg.TaskType.CAUSAL_LM)
    m = _tiny_gpt2()
    pm = pkg.get_peft_model(m, cfg)
    layers = [mod for mod in pm.modules() if isinstance(mod, pkg.lora.Linear)]
    assert len(layers) == 18 and all(l.fan_in_fan_out for l in layers)
    assert pm.get_nb_trainable_parameters()[0] == 405_504
    assert all(("lora_" in n) == p.requires_grad for n, p in pm.named_parameters())
    l0 = pm.base_model.model.transformer.h[0].attn.c_attn
    assert l0.scaling["default"] == 2.0 and l0.r["default"] == 8
    assert l0.lora_A["default"].weight.shape == (8, 768) and l0.lora_B["default"].weight.shape == (2304, 8)
    assert l0.lora_A["default"].weight.dtype == torch.float32
    assert float(l0.lora_B["default"].weight.abs().max()) == 0.0  # B = 0 at step 0
    bound = 1 / (768 ** 0.5)
    assert float(l0.lora_A["default"].weight.abs().max()) <= bound
    # attribute paths the reference relies on
    assert pm.base_model.model is m
    assert pm.lm_head is m.lm_head and pm.config is m.config
    sd = pm.state_dict()
    assert "base_model.model.transformer.h.0.attn.c_attn.base_layer.weight" in sd
    assert "base_model.model.transformer.h.0.attn.c_attn.lora_A.default.weight" in sd
    assert "base_model.model.transformer.h.5.mlp.c_proj.lora_B.default.weight" in sd


def test_opt_targets(pkg):
    """finetune_lora_opt_pp.py:104-112: six nn.Linear targets per decoder layer."""
    from transformers import OPTConfig, OPTForCausalLM

    m = OPTForCausalLM(OPTConfig(hidden_size=64, ffn_dim=256, num_hidden_layers=2, num_attention_heads=4,
                                 word_embed_proj_dim=64, vocab_size=128, max_position_embeddings=32))
    cfg = pkg.LoraConfig(r=8, lora_alpha=16, lora_dropout=0.05, bias="none",
                         target_modules=["q_proj", "k_proj", "v_proj", "out_proj", "fc1", "fc2"],
                         task_type=pkg.TaskType.CAUSAL_LM)
    pm = pkg.get_peft_model(m, cfg)
    layers = [mod for mod in pm.modules() if isinstance(mod, pkg.lora.Linear)]
    assert len(layers) == 12 and not any(l.fan_in_fan_out for l in layers)
    opt_model = pm.base_model.model.model  # finetune_lora_opt_pp.py:116
    assert len(opt_model.decoder.layers) == 2
    per_layer = 8 * (4 * (64 + 64) + (64 + 256) + (256 + 64))
    assert pm.get_nb_trainable_parameters()[0] == 2 * per_layer


def test_config_errors(pkg):
    with pytest.raises(NotImplementedError):
        pkg.LoraConfig(r=8, target_modules=["x"], bias="all")
    with pytest.raises(ValueError):
        pkg.LoraConfig(r=0, target_modules=["x"])
    with pytest.raises(ValueError, match="not found in the base model"):
        pkg.get_peft_model(nn.Sequential(nn.Linear(8, 8)), pkg.LoraConfig(r=4, target_modules=["nope"]))
    with pytest.raises(ValueError, match="unsupported LoRA target"):
        pkg.get_peft_model(nn.Sequential(nn.Conv2d(1, 1, 1)), pkg.LoraConfig(r=4, target_modules=["0"]))
    with pytest.raises(ValueError, match=r"\[1, 64\]"):
        pkg.get_peft_model(nn.Sequential(nn.Linear(8, 8)), pkg.LoraConfig(r=65, target_modules=["0"]))


def test_save_pretrained_layout_and_roundtrip(pkg, tmp_path):
    """finetune_lora_distilgpt2.py:167 / finetune_lora_opt_pp.py:239: adapter_config.json +
    adapter_model.safetensors with peft's key naming (adapter name stripped)."""
    from safetensors.torch import load_file

    cfg = pkg.LoraConfig(r=8, lora_alpha=16, target_modules=["c_attn", "c_proj"], lora_dropout=0.05,
                         task_type=pkg.TaskType.CAUSAL_LM)
    pm = pkg.get_peft_model(_tiny_gpt2(), cfg)
    with torch.no_grad():
        for n, p in pm.named_parameters():
            if "lora_B" in n:
                p.normal_()
    out = str(tmp_path / "adapter")
    pm.save_pretrained(out)
    meta = json.load(open(os.path.join(out, "adapter_config.json")))
    assert meta["peft_type"] == "LORA" and meta["r"] == 8 and meta["lora_alpha"] == 16
    assert meta["target_modules"] == ["c_attn", "c_proj"] and meta["task_type"] == "CAUSAL_LM"
    sd = load_file(os.path.join(out, "adapter_model.safetensors"))
    assert len(sd) == 36
    assert "base_model.model.transformer.h.0.attn.c_attn.lora_A.weight" in sd
    pm2 = pkg.PeftModel.from_pretrained(_tiny_gpt2(), out)
    a = pm.state_dict()["base_model.model.transformer.h.3.mlp.c_proj.lora_B.default.weight"]
    b = pm2.state_dict()["base_model.model.transformer.h.3.mlp.c_proj.lora_B.default.weight"]
    assert torch.equal(a, b)


def test_save_run_writes_adapter_and_run_meta(pkg, tmp_path):
    """finetune_lora_distilgpt2.py:162-185: rank-0 save of adapter (+ tokenizer) and run_meta.json."""
    pm = pkg.get_peft_model(_tiny_gpt2(), pkg.LoraConfig(r=8, target_modules=["c_attn"], task_type=pkg.TaskType.CAUSAL_LM))
    pkg.frozen.accelerate_lm_head(pm)  # the optional module surgery must not change what is saved
    meta = {"base_model": "distilgpt2", "world_size": 1, "seq_len": 128, "batch": 4, "accum": 32, "lr": 5e-5,
            "train_seconds": 1.5}
    out = str(tmp_path / "run")
    pkg.train.save_run(pm, out, meta)
    assert json.load(open(os.path.join(out, "run_meta.json"))) == meta
    assert os.path.exists(os.path.join(out, "adapter_config.json"))
    from safetensors.torch import load_file

    sd = load_file(os.path.join(out, "adapter_model.safetensors"))
    assert len(sd) == 12 and all(".lora_A.weight" in k or ".lora_B.weight" in k for k in sd)


def test_peft_shim_import(pkg):
    """`from peft import LoraConfig, get_peft_model, TaskType` (finetune_lora_distilgpt2.py:78)
    resolves to this library once <package>/peft_shim is on sys.path."""
    import importlib
    import sys

    shim = os.path.join(os.path.dirname(pkg.__file__), "peft_shim")
    sys.path.insert(0, shim)
    try:
        sys.modules.pop("peft", None)
        peft = importlib.import_module("peft")
        assert peft.LoraConfig is pkg.LoraConfig and peft.get_peft_model is pkg.get_peft_model
        assert peft.TaskType.CAUSAL_LM == "CAUSAL_LM" and peft.PeftModel is pkg.PeftModel
    finally:
        sys.path.remove(shim)
        sys.modules.pop("peft", None)


def test_merge_unmerge_math(pkg):
    lin = nn.Linear(16, 24)
    layer = pkg.lora.Linear(lin, r=4, lora_alpha=8)
    with torch.no_grad():
        layer.lora_B["default"].weight.normal_()
    w0 = lin.weight.detach().clone()
    delta = 2.0 * layer.lora_B["default"].weight @ layer.lora_A["default"].weight
    layer.merge()
    assert layer.merged and torch.allclose(lin.weight, w0 + delta, atol=1e-6)
    layer.unmerge()
    assert not layer.merged and torch.allclose(lin.weight, w0, atol=1e-6)
