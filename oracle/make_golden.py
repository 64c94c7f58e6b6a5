"""Generates tests/golden/lora_linear_golden.npz from the oracle (run from the repo root).

The reference has no fixtures for this path and peft cannot be imported here,
so these vectors are produced by oracle/lora_ref.py itself: they pin the oracle against drift
and give the GPU tests a file-based target that does not depend on re-running the oracle."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import lora_ref, philox  # noqa: E402

CASES = [  # M, K, N, r, w_is_KN, p, seed, offset
    (96, 256, 192, 8, 1, 0.05, 1234, 0),  # Conv1D layout (GPT-2 c_attn / c_proj family), dropout on
    (64, 192, 320, 8, 1, 0.0, 1, 4),  # Conv1D layout, eval-style (no dropout)
    (48, 320, 256, 8, 0, 0.05, 77, 8),  # nn.Linear layout (OPT q/k/v/out_proj, fc1, fc2 family)
    (40, 136, 72, 33, 0, 0.05, 5, 12),  # K, N not multiples of 64; r padded to 64 on the device
]


def main():
    out = {"n_cases": np.int64(len(CASES))}
    for i, (M, K, N, r, kn, p, seed, offset) in enumerate(CASES):
        g = torch.Generator().manual_seed(1000 + i)
        q = lambda t: t.to(torch.bfloat16).float()
        x = q(torch.randn(M, K, generator=g))
        W = q(torch.randn((K, N) if kn else (N, K), generator=g) * 0.02)
        b = q(torch.randn(N, generator=g) * 0.02)
        A = (torch.rand(r, K, generator=g) * 2 - 1) / np.sqrt(K)
        B = torch.randn(N, r, generator=g) * 0.02
        dy = q(torch.randn(M, N, generator=g))
        keep = torch.from_numpy(philox.keep_mask(M, K, p, seed, offset)) if p > 0 else None
        res = lora_ref.lora_linear_fwd_bwd(x, W, b, A, B, 16.0 / r, dy, keep, p, bool(kn))
        for k, v in dict(x=x, W=W, b=b, A=A, B=B, dy=dy, **{k: res[k] for k in ("y", "dx", "dA", "dB")}).items():
            out[f"c{i}_{k}"] = v.numpy()
        out[f"c{i}_meta"] = np.array([M, K, N, r, kn, round(p * 1e4), seed, offset], dtype=np.int64)
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                        "lora_linear_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
