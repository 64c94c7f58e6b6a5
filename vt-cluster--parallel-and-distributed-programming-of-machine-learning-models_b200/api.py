"""The slice of peft's public API that the reference uses, over the fused kernels.

    from peft import LoraConfig, get_peft_model, TaskType        finetune_lora_distilgpt2.py:78
    from peft import LoraConfig, get_peft_model, TaskType        finetune_lora_opt_pp.py:9
    from peft import PeftModel                                   finetune_lora_opt_pp.py:236
    peft_model.base_model.model.model                            finetune_lora_opt_pp.py:116
    peft_model.lm_head                                           finetune_lora_opt_pp.py:142
    trainer.model.save_pretrained(dir)                           finetune_lora_distilgpt2.py:167

Same names, argument meaning and error behaviour; the arithmetic is lora.Linear
in this package (sm_100a kernels), never peft's eager ops.
"""
from __future__ import annotations

import enum
import json
import os
import re
from dataclasses import asdict, dataclass, field
from typing import List, Optional, Union

import torch
import torch.nn as nn

from . import lora


class TaskType(str, enum.Enum):
    SEQ_CLS = "SEQ_CLS"
    SEQ_2_SEQ_LM = "SEQ_2_SEQ_LM"
    CAUSAL_LM = "CAUSAL_LM"
    TOKEN_CLS = "TOKEN_CLS"
    QUESTION_ANS = "QUESTION_ANS"
    FEATURE_EXTRACTION = "FEATURE_EXTRACTION"


class PeftType(str, enum.Enum):
    LORA = "LORA"


@dataclass
class LoraConfig:
    """peft.LoraConfig fields the reference sets (r, lora_alpha, target_modules, lora_dropout,
    bias, task_type) plus the defaults that shape the arithmetic."""

    r: int = 8
    lora_alpha: int = 8
    target_modules: Optional[Union[List[str], str]] = None
    lora_dropout: float = 0.0
    fan_in_fan_out: bool = False
    bias: str = "none"
    task_type: Optional[Union[str, TaskType]] = None
    init_lora_weights: bool = True
    use_rslora: bool = False
    modules_to_save: Optional[List[str]] = None
    inference_mode: bool = False
    base_model_name_or_path: Optional[str] = None
    peft_type: str = field(default="LORA")

    def __post_init__(self):
        if isinstance(self.task_type, str) and not isinstance(self.task_type, TaskType):
            self.task_type = TaskType(self.task_type)
        if isinstance(self.target_modules, (set, tuple)):
            self.target_modules = list(self.target_modules)
        if self.bias not in ("none", "all", "lora_only"):
            raise ValueError(f"Unknown bias setting {self.bias!r}")
        if self.bias != "none":
            raise NotImplementedError('only bias="none" is implemented (the reference never trains biases: '
                                      "finetune_lora_distilgpt2.py:101, finetune_lora_opt_pp.py:108)")
        if self.r <= 0:
            raise ValueError(f"`r` should be a positive integer value but the value passed is {self.r}")

    def to_dict(self):
        d = asdict(self)
        d["task_type"] = self.task_type.value if isinstance(self.task_type, TaskType) else self.task_type
        if isinstance(d.get("target_modules"), list):
            d["target_modules"] = sorted(d["target_modules"])
        return d


def _matches(key: str, target_modules) -> bool:
    if isinstance(target_modules, str):
        return re.fullmatch(target_modules, key) is not None
    return any(key == t or key.endswith("." + t) for t in target_modules)


class LoraModel(nn.Module):
    """peft's tuner: owns the (mutated) base model as `.model`."""

    prefix = "lora_"

    def __init__(self, model: nn.Module, config: LoraConfig, adapter_name: str = "default"):
        super().__init__()
        self.model = model
        self.peft_config = {adapter_name: config}
        self.active_adapter = adapter_name
        self._inject(config, adapter_name)
        self._mark_only_lora_as_trainable()

    def _inject(self, cfg: LoraConfig, adapter_name: str):
        if cfg.target_modules is None:
            raise ValueError("Please specify `target_modules` in `peft_config`")
        hits = [n for n, _ in self.model.named_modules() if n and _matches(n, cfg.target_modules)]
        if not hits:
            raise ValueError(f"Target modules {cfg.target_modules} not found in the base model. "
                             "Please check the target modules and try again.")
        for name in hits:
            parent_name, _, child = name.rpartition(".")
            parent = self.model.get_submodule(parent_name) if parent_name else self.model
            target = getattr(parent, child)
            if isinstance(target, lora.Linear):
                target.update_layer(adapter_name, cfg.r, cfg.lora_alpha, cfg.lora_dropout, cfg.init_lora_weights,
                                    cfg.use_rslora)
                continue
            new = lora.Linear(target, adapter_name, r=cfg.r, lora_alpha=cfg.lora_alpha,
                              lora_dropout=cfg.lora_dropout, fan_in_fan_out=cfg.fan_in_fan_out,
                              init_lora_weights=cfg.init_lora_weights, use_rslora=cfg.use_rslora)
            setattr(parent, child, new)

    def _mark_only_lora_as_trainable(self):
        for n, p in self.model.named_parameters():
            p.requires_grad_(self.prefix in n)

    def forward(self, *args, **kwargs):
        return self.model(*args, **kwargs)

    def __getattr__(self, name: str):
        try:
            return super().__getattr__(name)
        except AttributeError:
            if name == "model":
                raise
            return getattr(self.model, name)

    def lora_layers(self):
        return [m for m in self.model.modules() if isinstance(m, lora.Linear)]

    def merge_adapter(self):
        for m in self.lora_layers():
            m.merge()

    def unmerge_adapter(self):
        for m in self.lora_layers():
            m.unmerge()

    def enable_adapter_layers(self):
        for m in self.lora_layers():
            m.enable_adapters(True)

    def disable_adapter_layers(self):
        for m in self.lora_layers():
            m.enable_adapters(False)


class PeftModel(nn.Module):
    """peft.PeftModel / PeftModelForCausalLM: `.base_model` is the tuner, `.base_model.model` the
    original transformers model; unknown attributes forward down to it."""

    def __init__(self, model: nn.Module, peft_config: LoraConfig, adapter_name: str = "default"):
        super().__init__()
        self.active_adapter = adapter_name
        self.peft_type = PeftType.LORA
        self.base_model = LoraModel(model, peft_config, adapter_name)
        if getattr(model, "is_gradient_checkpointing", False) and hasattr(model, "enable_input_require_grads"):
            # peft does the same: with frozen embeddings, checkpointed blocks would otherwise see no
            # input that requires grad and skip the adapters' backward.
            model.enable_input_require_grads()

    @property
    def peft_config(self):
        return self.base_model.peft_config

    def forward(self, *args, **kwargs):
        return self.base_model(*args, **kwargs)

    def generate(self, *args, **kwargs):
        return self.base_model.model.generate(*args, **kwargs)

    def __getattr__(self, name: str):
        try:
            return super().__getattr__(name)
        except AttributeError:
            if name == "base_model":
                raise
            return getattr(self.base_model, name)

    def get_base_model(self):
        return self.base_model.model

    def get_nb_trainable_parameters(self):
        trainable = sum(p.numel() for p in self.parameters() if p.requires_grad)
        total = sum(p.numel() for p in self.parameters())
        return trainable, total

    def print_trainable_parameters(self):
        t, a = self.get_nb_trainable_parameters()
        print(f"trainable params: {t:,d} || all params: {a:,d} || trainable%: {100 * t / a:.4f}")

    # ---- adapter on-disk format (peft layout: adapter_config.json + adapter_model.safetensors) ----
    def adapter_state_dict(self, adapter_name: str = "default"):
        out = {}
        for k, v in self.state_dict().items():
            if "lora_" in k and f".{adapter_name}." in k:
                out[k.replace(f".{adapter_name}.", ".")] = v.detach().cpu().contiguous()
        return out

    def save_pretrained(self, save_directory: str, safe_serialization: bool = True, **_):
        if os.path.isfile(save_directory):
            raise ValueError(f"Provided path ({save_directory}) should be a directory, not a file")
        os.makedirs(save_directory, exist_ok=True)
        sd = self.adapter_state_dict(self.active_adapter)
        if safe_serialization:
            from safetensors.torch import save_file

            save_file(sd, os.path.join(save_directory, "adapter_model.safetensors"), metadata={"format": "pt"})
        else:
            torch.save(sd, os.path.join(save_directory, "adapter_model.bin"))
        cfg = self.peft_config[self.active_adapter].to_dict()
        cfg["inference_mode"] = True
        if cfg.get("base_model_name_or_path") is None:
            cfg["base_model_name_or_path"] = getattr(getattr(self.base_model.model, "config", None), "_name_or_path",
                                                     None)
        with open(os.path.join(save_directory, "adapter_config.json"), "w") as f:
            json.dump(cfg, f, indent=2, sort_keys=True)

    def load_adapter_state(self, save_directory: str, adapter_name: str = "default"):
        path = os.path.join(save_directory, "adapter_model.safetensors")
        if os.path.exists(path):
            from safetensors.torch import load_file

            sd = load_file(path)
        else:
            sd = torch.load(os.path.join(save_directory, "adapter_model.bin"), map_location="cpu")
        own = self.state_dict()
        for k, v in sd.items():
            kk = re.sub(r"\.(lora_[AB])\.weight$", rf".\1.{adapter_name}.weight", k)
            if kk not in own:
                raise KeyError(f"unexpected adapter key {k}")
            own[kk].copy_(v)

    @classmethod
    def from_pretrained(cls, model: nn.Module, model_id: str, adapter_name: str = "default", is_trainable=False, **_):
        with open(os.path.join(model_id, "adapter_config.json")) as f:
            raw = json.load(f)
        known = {k: raw[k] for k in LoraConfig.__dataclass_fields__ if k in raw}
        cfg = LoraConfig(**known)
        cfg.inference_mode = not is_trainable
        pm = cls(model, cfg, adapter_name)
        pm.load_adapter_state(model_id, adapter_name)
        if not is_trainable:
            for p in pm.parameters():
                p.requires_grad_(False)
        return pm


PeftModelForCausalLM = PeftModel


def get_peft_model(model: nn.Module, peft_config: LoraConfig, adapter_name: str = "default") -> PeftModel:
    """peft.get_peft_model: wrap every target module, freeze the rest, return the PeftModel."""
    if not isinstance(peft_config, LoraConfig):
        raise TypeError("peft_config must be a LoraConfig (only the LoRA tuner exists in this library)")
    if peft_config.base_model_name_or_path is None:
        peft_config.base_model_name_or_path = getattr(getattr(model, "config", None), "_name_or_path", None)
    return PeftModel(model, peft_config, adapter_name)
