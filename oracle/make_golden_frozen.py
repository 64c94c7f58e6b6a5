"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/frozen_golden.npz (run from the repo root).

Vectors for the callers either side of the LoRA path, produced on the CPU in fp32 by the
code the reference's scripts reach through transformers / torch:
  * frozen GPT-2 `mlp.c_fc` (transformers Conv1D, weight [K, N]) followed by NewGELUActivation, and a frozen
    nn.Linear followed by ReLU (OPT fc1): y and dX for a given dY;
  * the causal-LM loss of DataCollatorForLanguageModeling labels (-100 = ignored; finetune_lora_distilgpt2.py:114,
    finetune_lora_opt_pp.py:143-153): shifted F.cross_entropy and its gradient with respect to the logits.
Inputs are rounded to bf16 first so that the 16-bit device kernels see exactly these values."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import lora_ref  # noqa: E402


def q(t):
    return t.to(torch.bfloat16).float()


def main():
    from transformers.activations import NewGELUActivation

    out = {}
    g = torch.Generator().manual_seed(2024)
    for name, (M, K, N, act, kn) in {"gelu": (72, 128, 264, "gelu_new", 1), "relu": (40, 136, 128, "relu", 0)}.items():
        x = q(torch.randn(M, K, generator=g)).requires_grad_(True)
        W = q(torch.randn((K, N) if kn else (N, K), generator=g) * 0.05)
        b = q(torch.randn(N, generator=g) * 0.1)
        dy = q(torch.randn(M, N, generator=g))
        z = torch.addmm(b, x, W) if kn else torch.nn.functional.linear(x, W, b)  # Conv1D.forward / nn.Linear.forward
        y = NewGELUActivation()(z) if act == "gelu_new" else torch.relu(z)
        y.backward(dy)
        for k, v in dict(x=x.detach(), W=W, b=b, dy=dy, y=y.detach(), dx=x.grad).items():
            out[f"{name}_{k}"] = v.numpy()
        out[f"{name}_meta"] = np.array([M, K, N, kn], dtype=np.int64)
    B, S, V = 3, 21, 1003
    logits = q(torch.randn(B, S, V, generator=g) * 2).requires_grad_(True)
    labels = torch.randint(0, V, (B, S), generator=g)
    labels[1, 4:9] = -100
    loss = lora_ref.causal_lm_loss(logits, labels)
    loss.backward()
    out["ce_logits"], out["ce_labels"] = logits.detach().numpy(), labels.numpy()
    out["ce_loss"], out["ce_dlogits"] = np.float64(loss.item()), logits.grad.numpy()
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "frozen_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
