"""Drop-in `peft` for the reference's training scripts.

With `<package>/peft_shim` first on PYTHONPATH, the reference's own import lines
    from peft import LoraConfig, get_peft_model, TaskType   (finetune_lora_distilgpt2.py:78,
                                                              finetune_lora_opt_pp.py:9)
    from peft import PeftModel                               (finetune_lora_opt_pp.py:236)
resolve to the B200 implementation without editing the scripts.
"""
import importlib.util
import os
import sys

_PKG_DIR = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
_ALIAS = "lora_b200"

if _ALIAS not in sys.modules:
    _spec = importlib.util.spec_from_file_location(
        _ALIAS, os.path.join(_PKG_DIR, "__init__.py"), submodule_search_locations=[_PKG_DIR]
    )
    _mod = importlib.util.module_from_spec(_spec)
    sys.modules[_ALIAS] = _mod
    _spec.loader.exec_module(_mod)
_impl = sys.modules[_ALIAS]

LoraConfig = _impl.LoraConfig
TaskType = _impl.TaskType
PeftType = _impl.PeftType
PeftModel = _impl.PeftModel
PeftModelForCausalLM = _impl.PeftModelForCausalLM
LoraModel = _impl.LoraModel
get_peft_model = _impl.get_peft_model

__version__ = "0.0.0+lora_b200"
__all__ = ["LoraConfig", "TaskType", "PeftType", "PeftModel", "PeftModelForCausalLM", "LoraModel", "get_peft_model"]
