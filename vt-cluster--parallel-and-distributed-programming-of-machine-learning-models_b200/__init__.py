"""B200-native fused LoRA-linear path (sm_100a) behind peft's LoraConfig / get_peft_model API.

The directory name is not a Python identifier; load it through
`__graft_entry__.load_package()` (alias `lora_b200`) or put `<this dir>/peft_shim`
on PYTHONPATH so that the reference's `from peft import ...` lines resolve here.
"""
from . import binding  # noqa: F401  (ctypes C-ABI; fails loudly at first use if the .so is missing)
from . import lora  # noqa: F401
from . import train  # noqa: F401
from . import pipeline  # noqa: F401
from . import frozen  # noqa: F401
from . import comm  # noqa: F401
from .api import (  # noqa: F401
    LoraConfig,
    LoraModel,
    PeftModel,
    PeftModelForCausalLM,
    PeftType,
    TaskType,
    get_peft_model,
)
from .lora import fused_lora_linear  # noqa: F401
from ._build import build  # noqa: F401

__all__ = [
    "LoraConfig", "LoraModel", "PeftModel", "PeftModelForCausalLM", "PeftType", "TaskType", "get_peft_model",
    "fused_lora_linear", "binding", "lora", "train", "pipeline", "frozen", "comm", "build",
]
