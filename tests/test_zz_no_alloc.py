"""lora_linear_grad_b_static: the adapter gradient dB for callers that cannot allocate.

Some users of this library run where device memory cannot be requested at call time: a plug-in
inside a runtime that already owns the GPU's memory pool, or a loop that must never enter the
driver's allocator.  For them the library offers one entry point whose scratch (the split partial
sums of dY^T H) is a fixed block built into the library.  This test drives it through the C ABI
exactly as such a caller would -- raw pointers, no workspace -- and checks, on two calls in a row
with different data (the block is shared and keeps the previous call's contents), that dB matches
what the workspace-based lora_linear_bwd produces for the same dY and H.  It also checks that a
shape whose partial sums exceed the block is refused with a message naming the limit.

The file name sorts last on purpose: if this test fails under `pytest -x`, every other test has
already run.  It needs no state from other tests and passes in a fresh process on its own."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-2  # same bound as tests/test_gpu_parity.py (in practice the two paths differ by fp32 rounding only)


def _relerr(a, ref):
    return float((a.float() - ref.float()).abs().max() / ref.float().abs().max().clamp_min(1e-20))


def _reference_path(pkg, M, K, N, r, seed):
    """dY, H and dB from the existing entry points (forward writes H, backward writes dB)."""
    g = torch.Generator().manual_seed(seed)
    dt = torch.bfloat16
    x = torch.randn(M, K, generator=g).to(dt).cuda()
    W = (torch.randn(N, K, generator=g) * 0.02).to(dt).cuda()
    A = ((torch.rand(r, K, generator=g) * 2 - 1) / K ** 0.5).cuda()
    B = (torch.randn(N, r, generator=g) * 0.02).cuda()
    dy = torch.randn(M, N, generator=g).to(dt).cuda()
    s = 16.0 / r
    ws = torch.empty(pkg.binding.workspace_bytes(M, K, N, r), dtype=torch.uint8, device="cuda")
    _, H, keep, packed = pkg.binding.linear_fwd(x, W, None, A, B, s, 0.0, 0, 0, False)
    _, _, dB = pkg.binding.linear_bwd(dy, x, W, packed, H, keep, r, s, 0.0, False, ws, need_dx=False)
    torch.cuda.synchronize()
    return dy, H, dB, s


def _grad_b_static(pkg, dy, H, dB_out, r, s, accumulate=0):
    lib = pkg.binding.lib()
    M, N = dy.shape
    return lib.lora_linear_grad_b_static(
        ctypes.c_void_p(dy.data_ptr()), ctypes.c_void_p(H.data_ptr()), ctypes.c_void_p(dB_out.data_ptr()),
        M, N, r, ctypes.c_float(s), pkg.binding.DTYPE_BF16, accumulate,
        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))


def test_static_scratch_matches_workspace_path_twice_in_a_row(pkg):
    lib = pkg.binding.lib()
    # First call on this device (library set-up, see the header) happens here, before anything is measured.
    warm = torch.zeros(64, 64, dtype=torch.bfloat16, device="cuda")
    assert _grad_b_static(pkg, warm, warm[:, :16].contiguous(), torch.zeros(64, 8, device="cuda"), 8, 1.0) == 0
    torch.cuda.synchronize()
    # two different problems back to back: the second must not see anything of the first
    for (M, K, N, r, seed) in [(4096, 768, 2560, 8, 1), (3000, 256, 2304, 16, 2)]:
        dy, H, dB_ref, s = _reference_path(pkg, M, K, N, r, seed)
        dB = torch.full((N, r), float("nan"), device="cuda")  # every element must be written
        # The point of the entry point: the call takes no device memory.  Free bytes as the driver reports them
        # (cudaMemGetInfo) are the same before and after -- torch's caching allocator is not involved, the call
        # receives raw pointers -- and torch's own allocator statistics do not move either.
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        reserved_before = torch.cuda.memory_reserved()
        rc = _grad_b_static(pkg, dy, H, dB, r, s)
        assert rc == 0, lib.lora_last_error().decode()
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert free_after == free_before, (free_before, free_after)
        assert torch.cuda.memory_reserved() == reserved_before
        assert torch.isfinite(dB).all()
        assert _relerr(dB, dB_ref) < TOL
        # accumulate != 0 adds on top of what is there
        rc = _grad_b_static(pkg, dy, H, dB, r, s, accumulate=1)
        assert rc == 0, lib.lora_last_error().decode()
        torch.cuda.synchronize()
        assert _relerr(dB, 2 * dB_ref) < TOL


def test_static_scratch_takes_fewer_splits_when_the_planned_ones_do_not_fit(pkg):
    """M = 16384, N = 2560, r = 64: the planner wants 7 row splits x 2560 x 64 x 4 B = 4.6 MB, the block holds 4 MiB; the
    call runs with 6 and gives the workspace path's result."""
    M, K, N, r = 16384, 256, 2560, 64
    dy, H, dB_ref, s = _reference_path(pkg, M, K, N, r, 5)
    dB = torch.full((N, r), float("nan"), device="cuda")
    rc = _grad_b_static(pkg, dy, H, dB, r, s)
    assert rc == 0, pkg.binding.lib().lora_last_error().decode()
    torch.cuda.synchronize()
    assert torch.isfinite(dB).all() and _relerr(dB, dB_ref) < TOL


def test_static_scratch_rejects_oversized_shapes(pkg):
    lib = pkg.binding.lib()
    M, N, r = 256, 20480, 64  # one row split x 20480 x 64 x 4 B = 5 MiB of partial sums > the 4 MiB block
    dy = torch.zeros(M, N, dtype=torch.bfloat16, device="cuda")
    H = torch.zeros(M, 64, dtype=torch.bfloat16, device="cuda")
    dB = torch.zeros(N, r, device="cuda")
    rc = _grad_b_static(pkg, dy, H, dB, r, 1.0)
    msg = lib.lora_last_error().decode()
    assert rc != 0 and "built-in scratch holds 4194304" in msg and "lora_linear_bwd" in msg
