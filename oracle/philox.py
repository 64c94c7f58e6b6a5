"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the dropout-mask generator.

Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as
1, 2, 3", SC'11), consumed exactly as csrc/common.cuh::dropout_keep_bits does:
one call per aligned group of 8 consecutive elements of the [M, K] adapter
input, counter = (g_lo, g_hi, offset_lo, offset_hi), key = (seed_lo, seed_hi),
element e of the group takes the low 14 bits of the e-th 16-bit lane and
survives iff lane14 >= round(p * 16384).

The reference draws its mask from torch's CPU generator inside peft's
nn.Dropout (lora_dropout=0.05 at finetune_lora_distilgpt2.py:101 and
finetune_lora_opt_pp.py:107); that stream cannot be reproduced on a GPU, so
parity tests feed THIS mask to both sides ("Dropout
parity and replay").  Known-answer pins for the generator itself are the
Random123 vectors in tests/test_oracle.py.
"""
from __future__ import annotations

import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10.  Counters are uint32 arrays (broadcastable), keys python ints."""
    c0 = np.asarray(c0, dtype=np.uint64)
    c1 = np.asarray(c1, dtype=np.uint64)
    c2 = np.asarray(c2, dtype=np.uint64)
    c3 = np.asarray(c3, dtype=np.uint64)
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c0
        p1 = _M1 * c2
        n0 = (p1 >> np.uint64(32)) ^ c1 ^ np.uint64(k0)
        n1 = p1 & _MASK32
        n2 = (p0 >> np.uint64(32)) ^ c3 ^ np.uint64(k1)
        n3 = p0 & _MASK32
        c0, c1, c2, c3 = n0, n1, n2, n3
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return (c0.astype(np.uint32), c1.astype(np.uint32), c2.astype(np.uint32), c3.astype(np.uint32))


def thresh14(p: float) -> int:
    """Same rounding as the host code in csrc/lora_b200.cu (float32 p * 16384 + 0.5, truncated)."""
    return int(np.float32(p) * np.float32(16384.0) + np.float32(0.5))


def keep_bytes(M: int, K: int, p: float, seed: int, offset: int) -> np.ndarray:
    """Packed keep-bits, shape [M, K//8] uint8; bit e of byte (m, k//8) is element (m, 8*(k//8)+e)."""
    assert K % 8 == 0
    groups = np.arange(M * (K // 8), dtype=np.uint64)
    g_lo = (groups & _MASK32).astype(np.uint32)
    g_hi = (groups >> np.uint64(32)).astype(np.uint32)
    o_lo = np.uint32(offset & 0xFFFFFFFF)
    o_hi = np.uint32((offset >> 32) & 0xFFFFFFFF)
    r = philox4x32_10(g_lo, g_hi, o_lo, o_hi, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    t = thresh14(p)
    bits = np.zeros(groups.shape, dtype=np.uint8)
    for e in range(8):
        lane = (r[e >> 1] >> np.uint32((e & 1) * 16)) & np.uint32(0x3FFF)
        bits |= ((lane >= t).astype(np.uint8) << np.uint8(e))
    return bits.reshape(M, K // 8)


def keep_mask(M: int, K: int, p: float, seed: int, offset: int) -> np.ndarray:
    """Boolean keep mask [M, K] (True = element survives dropout)."""
    kb = keep_bytes(M, K, p, seed, offset)
    return np.unpackbits(kb, axis=1, bitorder="little").astype(bool).reshape(M, K)
