"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the dropout-mask generator.

Philox4x32-10 (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as
1, 2, 3", SC'11), consumed exactly as csrc/common.cuh ("layout of the dropout
stream") does: the [M, K] adapter input is cut into 16x16 element blocks; block
(R, C) takes 32 calls with counter = (g_lo, g_hi, offset_lo, offset_hi),
g = (R * ceil(K/16) + C) * 32 + t, key = (seed_lo, seed_hi); element (i, j) of
the block belongs to call t = 4*(i%8) + (j%8)//2 and takes the low 14 bits of
16-bit lane e = 4*(j//8) + 2*(i//8) + j%2 of the call's 128 bits (call t draws
for the eight elements lane t of a warp holds in the A fragment of an m16n8k16
MMA over the block); it survives iff lane14 >= round(p * 16384).

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


def _keep_block_rows(block_rows: np.ndarray, K: int, p: float, seed: int, offset: int) -> np.ndarray:
    """Keep decisions of whole 16-row block rows: bool [len(block_rows), 16, 16 * ceil(K/16)]."""
    C = (K + 15) // 16
    block_rows = np.asarray(block_rows, dtype=np.uint64)
    blk = (block_rows[:, None] * np.uint64(C) + np.arange(C, dtype=np.uint64)[None, :]).reshape(-1)  # [nR*C]
    g = blk[:, None] * np.uint64(32) + np.arange(32, dtype=np.uint64)[None, :]  # [nR*C, 32]
    g_lo = (g & _MASK32).astype(np.uint32)
    g_hi = (g >> np.uint64(32)).astype(np.uint32)
    o_lo = np.uint32(offset & 0xFFFFFFFF)
    o_hi = np.uint32((offset >> 32) & 0xFFFFFFFF)
    r = philox4x32_10(g_lo, g_hi, o_lo, o_hi, seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    t14 = thresh14(p)
    # lanes[block, t, e]: e = 2*w + h is half h of 32-bit word w of call t
    lanes = np.empty((blk.shape[0], 32, 8), dtype=np.uint32)
    for w in range(4):
        lanes[:, :, 2 * w] = r[w] & np.uint32(0x3FFF)
        lanes[:, :, 2 * w + 1] = (r[w] >> np.uint32(16)) & np.uint32(0x3FFF)
    keep_bte = lanes >= t14
    i = np.arange(16)[:, None]
    j = np.arange(16)[None, :]
    t_of = 4 * (i % 8) + (j % 8) // 2  # [16, 16]
    e_of = 4 * (j // 8) + 2 * (i // 8) + j % 2
    blocks = keep_bte[:, t_of, e_of]  # [nR*C, 16, 16]
    return blocks.reshape(len(block_rows), C, 16, 16).transpose(0, 2, 1, 3).reshape(len(block_rows), 16, C * 16)


def keep_mask(M: int, K: int, p: float, seed: int, offset: int) -> np.ndarray:
    """Boolean keep mask [M, K] (True = element survives dropout)."""
    assert K % 8 == 0
    R = (M + 15) // 16
    full = _keep_block_rows(np.arange(R), K, p, seed, offset).reshape(R * 16, -1)
    return np.ascontiguousarray(full[:M, :K])


def keep_rows(rows, K: int, p: float, seed: int, offset: int) -> np.ndarray:
    """Rows `rows` of keep_mask(M, K, ...) for any M > max(rows), without drawing the whole mask."""
    rows = np.asarray(rows)
    blocks = _keep_block_rows(rows // 16, K, p, seed, offset)  # [n, 16, C*16]
    return np.ascontiguousarray(blocks[np.arange(len(rows)), rows % 16, :K])


def keep_bytes(M: int, K: int, p: float, seed: int, offset: int) -> np.ndarray:
    """Packed keep-bits, shape [M, K//8] uint8; bit e of byte (m, k//8) is element (m, 8*(k//8)+e)."""
    return np.packbits(keep_mask(M, K, p, seed, offset), axis=1, bitorder="little")
