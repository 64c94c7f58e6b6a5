"""CPU-side checks of the drop-in boundary: the shared library loads without a GPU and exports
every symbol include/lora_b200.h declares; pure host-side entry points behave; GPU entry points
fail loudly (no CPU fallback) instead of computing something else."""
import os
import re

import pytest
import torch

REPO_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(REPO_ROOT, "include", "lora_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(lora_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.binding.lib()
    names = _declared()
    assert names == pkg.binding.exported_symbols()
    for n in names:
        assert getattr(lib, n) is not None


def test_host_only_entry_points(pkg):
    b = pkg.binding
    assert [b.rpad(r) for r in (1, 8, 16, 17, 32, 33, 64)] == [16, 16, 16, 32, 32, 64, 64]
    w1 = b.workspace_bytes(16384, 2560, 2560, 8)
    w2 = b.workspace_bytes(16384, 2560, 2560, 64)
    assert 0 < w1 < w2
    assert b.workspace_bytes(0, 8, 8, 8) == 0 and b.workspace_bytes(8, 8, 8, 65) == 0
    # packed adapter operands: four 16-bit copies of [rp, K] / [N, rp], each padded to 256 bytes
    assert b.packed_bytes(2560, 2560, 8) == 4 * 16 * 2560 * 2 and b.packed_bytes(776, 520, 33) >= 2 * 64 * (776 + 520) * 2
    assert b.packed_bytes(0, 8, 8) == 0
    assert b.lib().lora_last_error() == b""


def test_workspace_plan_is_one_wave(pkg):
    """The backward's scratch = dH [M, rp] (16-bit) + two counter tables + the split partial sums of dB and dA.  The
    column projection plans ONE wave of at most 148 CTAs (lora_b200.cu, colproj_plan), so the partial sums can never
    exceed 148 column-tile-splits of 128 columns x rp fp32 (plus what a single split needs when there are more column
    tiles than SMs) -- a host-only property of the planner, checked over the sweep's shapes and some ragged ones."""
    b = pkg.binding
    for (M, K, N, r) in [(16384, 2560, 2560, 8), (16384, 2560, 10240, 64), (4096, 10240, 2560, 16), (65536, 768, 2304, 64),
                         (130, 776, 520, 33), (1, 8, 8, 1), (2048, 2560, 2560, 8), (256, 768, 19456, 8), (300, 3072, 768, 24)]:
        rp = b.rpad(r)
        w = b.workspace_bytes(M, K, N, r)
        fixed = ((M * rp * 2 + 255) // 256) * 256 + 2 * 4096 * 4
        tiles = -(-K // 128) + -(-N // 128)
        one_split = (K + N) * rp * 4
        wave = 148 * 128 * rp * 4
        assert fixed + one_split <= w <= fixed + max(one_split, wave) + (K + N) * rp * 4 + 1024, (M, K, N, r, w)
        if tiles > 148:
            assert w <= fixed + one_split + 1024


def test_no_cpu_fallback(pkg):
    x = torch.zeros(8, 16, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.binding.linear_base(x, x, None, False)
    lin = torch.nn.Linear(16, 16).to(torch.bfloat16)
    layer = pkg.lora.Linear(lin, r=4, lora_alpha=8)
    with pytest.raises(RuntimeError):
        layer(x)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU error path")
def test_gpu_entry_point_errors_without_device(pkg):
    lib = pkg.binding.lib()
    rc = lib.lora_linear_base(1 << 12, 1 << 12, None, 1 << 12, 8, 8, 8, 0, 0, None)
    assert rc != 0 and len(lib.lora_last_error()) > 0
