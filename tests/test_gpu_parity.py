"""GPU parity: every entry point of include/lora_b200.h against the CPU oracle (oracle/).

Tolerance (BASELINE.json north_star): fused-layer outputs and adapter gradients within 2e-2
max relative error in bf16/fp16 against the fp32 CPU oracle fed the same 16-bit-rounded inputs
and the SAME dropout mask (oracle/philox.py restates the device generator).  Two numbers are
checked against that bound, both 2e-2:
  * relerr     = max|gpu - ref| / max|ref|                      (error relative to the tensor's scale)
  * relerr_vec = the same ratio taken separately for every vector along the tensor's long dimension -- every
                 token row of Y / H / dX, every rank row of dA [r,K], every rank column of dB [N,r] -- worst
                 vector reported: no row may hide behind a larger one.  (A bound relative to each single
                 ELEMENT is not meaningful for sums that pass through 16-bit H / dH: elements that cancel
                 to ~0 carry the absolute rounding error of their row.)
Integer/bit work (the keep-bits) must match exactly.
"""
import os

import numpy as np
import pytest
import torch

from oracle import lora_ref, philox

pytestmark = pytest.mark.gpu

TOL = 2e-2
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "lora_linear_golden.npz")


def relerr(a, ref):
    a = a.detach().float().cpu()
    ref = ref.detach().float().cpu()
    return float((a - ref).abs().max() / ref.abs().max().clamp_min(1e-20))


def relerr_vec(a, ref):
    a = a.detach().float().cpu()
    ref = ref.detach().float().cpu()
    dim = 1 if ref.shape[1] >= ref.shape[0] or ref.shape[1] > 64 else 0  # reduce over the long dimension
    num = (a - ref).abs().amax(dim=dim)
    den = ref.abs().amax(dim=dim)
    ok = den > 0
    if not bool(ok.any()):
        return float(num.max())
    return float((num[ok] / den[ok]).max()) if bool((num[~ok] == 0).all()) else float("inf")


def make_case(M, K, N, r, w_is_KN, dtype, bias=True, seed=0, b_zero=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(M, K, generator=g)
    W = torch.randn((K, N) if w_is_KN else (N, K), generator=g) * 0.02
    b = torch.randn(N, generator=g) * 0.02 if bias else None
    bound = 1.0 / np.sqrt(K)  # kaiming_uniform(a=sqrt(5)) on [r, K]
    A = (torch.rand(r, K, generator=g) * 2 - 1) * bound
    B = torch.zeros(N, r) if b_zero else torch.randn(N, r, generator=g) * 0.02
    dy = torch.randn(M, N, generator=g)
    q = lambda t: None if t is None else t.to(dtype).float()  # what the 16-bit kernels actually see
    return dict(x=q(x), W=q(W), b=q(b), A=A, B=B, dy=q(dy))


def to_dev(c, dtype):
    d = {}
    for k, v in c.items():
        if v is None:
            d[k] = None
        elif k in ("A", "B"):
            d[k] = v.cuda()
        else:
            d[k] = v.to(dtype).cuda()
    return d


FWD_BWD_CASES = [
    # M,    K,     N,    r,  p,    w_is_KN, dtype,          bias
    (512, 768, 2304, 8, 0.0, True, torch.bfloat16, True),  # distilgpt2 attn.c_attn (BASELINE cfg1 M)
    (512, 768, 768, 8, 0.05, True, torch.bfloat16, True),  # distilgpt2 attn.c_proj
    (300, 3072, 768, 8, 0.05, True, torch.bfloat16, True),  # distilgpt2 mlp.c_proj, ragged M
    (1024, 2560, 2560, 8, 0.05, False, torch.bfloat16, True),  # OPT q/k/v/out_proj
    (256, 2560, 10240, 16, 0.0, False, torch.bfloat16, True),  # OPT fc1, r=16
    (256, 10240, 2560, 64, 0.05, False, torch.bfloat16, True),  # OPT fc2, r=64 (wide-rank epilogue)
    (384, 2560, 2560, 8, 0.05, False, torch.float16, True),  # fp16 (BASELINE cfg5)
    (384, 768, 2304, 8, 0.05, True, torch.float16, False),  # fp16 Conv1D, no bias
    (1, 768, 768, 4, 0.05, False, torch.bfloat16, True),  # single token, r < 8
    (130, 776, 520, 33, 0.05, False, torch.bfloat16, True),  # K, N tails (not multiples of 64), r -> r_pad 64
    (130, 776, 520, 24, 0.0, True, torch.bfloat16, False),  # same tails, Conv1D, r -> r_pad 32
    # the flavour every adapter of the reference takes (r = 8, p = 0.05: CUDA-core masked epilogue) on ragged
    # K / N: keep-bit tail loads, partial 32-column chunks, K not a multiple of 16 (half dropout blocks)
    (130, 776, 520, 8, 0.05, False, torch.bfloat16, True),
    (130, 776, 520, 8, 0.05, True, torch.bfloat16, True),  # its Conv1D twin
    (257, 520, 776, 16, 0.05, False, torch.float16, False),  # fp16, r = 16, M one past a 256-row pair tile
    (200, 2560, 2560, 24, 0.05, False, torch.bfloat16, True),  # r -> r_pad 32 masked epilogue (smem slice flavour)
    # rank 64 / 32 over many row tiles and several row splits of the column projections (more tiles than ring
    # slots, split fold by the last CTA) and a contraction >= 2048 in the backward: the two-pass masked dX
    (2400, 2560, 2560, 64, 0.05, False, torch.bfloat16, True),
    (1300, 768, 2304, 32, 0.05, True, torch.float16, True),
    # more column tiles than SMs in the column projection (152 + 6 tiles of 128: one split each, two waves)
    (256, 768, 19456, 8, 0.05, False, torch.bfloat16, False),
    # row projections: up to 4736 rows the one-slab, eight-warp kernel (every case above), beyond it the four-slab one
    # (these two, ragged M, fewer CTAs than SMs; the full-size tests have more)
    (5000, 768, 768, 8, 0.05, True, torch.bfloat16, True),
    (4800, 2560, 2560, 64, 0.05, False, torch.float16, False),
]


@pytest.mark.parametrize("M,K,p,seed,offset", [(64, 768, 0.05, 1234, 0), (300, 2560, 0.1, 2**40 + 7, 2**33 + 4),
                                               (1, 8, 0.5, 0, 0), (1000, 776, 0.05, 99, 12)])
def test_keep_mask_bit_exact(pkg, M, K, p, seed, offset):
    got = pkg.binding.dropout_keep_mask(M, K, p, seed, offset).cpu().numpy()
    ref = philox.keep_bytes(M, K, p, seed, offset)
    assert np.array_equal(got, ref)


@pytest.mark.parametrize("M,K,N,w_is_KN,dtype", [
    (256, 768, 768, False, torch.bfloat16),
    (256, 768, 768, True, torch.bfloat16),
    (1000, 2560, 2560, False, torch.bfloat16),
    (1000, 2560, 2560, True, torch.float16),
    (4096, 3072, 768, True, torch.bfloat16),
    (130, 776, 520, False, torch.bfloat16),
    (130, 776, 520, True, torch.bfloat16),
])
def test_base_gemm(pkg, M, K, N, w_is_KN, dtype):
    c = make_case(M, K, N, 8, w_is_KN, dtype)
    d = to_dev(c, dtype)
    y = pkg.binding.linear_base(d["x"], d["W"], d["b"], w_is_KN)
    torch.cuda.synchronize()
    ref = torch.addmm(c["b"], c["x"], c["W"]) if w_is_KN else torch.nn.functional.linear(c["x"], c["W"], c["b"])
    assert relerr(y, ref) < TOL


def _run_fwd_bwd(pkg, M, K, N, r, p, w_is_KN, dtype, bias, b_zero=False, seed=11, offset=4, case=None):
    c = case if case is not None else make_case(M, K, N, r, w_is_KN, dtype, bias, b_zero=b_zero)
    d = to_dev(c, dtype)
    s = 16.0 / r
    ws = torch.empty(pkg.binding.workspace_bytes(M, K, N, r), dtype=torch.uint8, device="cuda")
    Y, H, keep, packed = pkg.binding.linear_fwd(d["x"], d["W"], d["b"], d["A"], d["B"], s, p, seed, offset, w_is_KN)
    dX, dA, dB = pkg.binding.linear_bwd(d["dy"], d["x"], d["W"], packed, H, keep, r, s, p, w_is_KN, ws)
    torch.cuda.synchronize()
    # a backward from separately packed operands (lora_linear_pack) must give bit-identical adapter gradients
    packed2 = pkg.binding.pack_adapters(d["A"], d["B"], s, dtype)
    assert torch.equal(packed2, packed)
    _, dA2, dB2 = pkg.binding.linear_bwd(d["dy"], d["x"], d["W"], packed2, H, keep, r, s, p, w_is_KN, ws, need_dx=False)
    assert torch.equal(dA2, dA) and torch.equal(dB2, dB)
    keep_ref = torch.from_numpy(philox.keep_mask(M, K, p, seed, offset)) if p > 0 else None
    if p > 0:
        assert np.array_equal(keep.cpu().numpy(), philox.keep_bytes(M, K, p, seed, offset))
    else:
        assert keep is None
    ref = lora_ref.lora_linear_fwd_bwd(c["x"], c["W"], c["b"], c["A"], c["B"], s, c["dy"], keep_ref, p, w_is_KN)
    return dict(Y=Y, H=H, dX=dX, dA=dA, dB=dB, ws=ws, d=d, c=c, s=s, keep=keep, packed=packed), ref


def _all_errs(out, ref, r):
    pairs = {"y": (out["Y"], ref["y"]), "h": (out["H"][:, :r], ref["h"]), "dx": (out["dX"], ref["dx"]),
             "dA": (out["dA"], ref["dA"]), "dB": (out["dB"], ref["dB"])}
    errs = {k: relerr(a, b) for k, (a, b) in pairs.items()}
    errs.update({k + "_vec": relerr_vec(a, b) for k, (a, b) in pairs.items()})
    return errs


@pytest.mark.parametrize("M,K,N,r,p,w_is_KN,dtype,bias", FWD_BWD_CASES)
def test_fwd_bwd_parity(pkg, M, K, N, r, p, w_is_KN, dtype, bias):
    out, ref = _run_fwd_bwd(pkg, M, K, N, r, p, w_is_KN, dtype, bias)
    errs = _all_errs(out, ref, r)
    assert all(np.isfinite(v) and v < TOL for v in errs.values()), errs
    if r < out["H"].shape[1]:
        assert float(out["H"][:, r:].float().abs().max()) == 0.0  # padding columns stay zero


def test_golden_fixture_through_the_gpu(pkg):
    """tests/golden/lora_linear_golden.npz (oracle/make_golden.py) fed to the device path: the file, not a fresh
    oracle run, is the target -- inputs, dropout (seed, offset) and expected y / dx / dA / dB all come from it."""
    g = np.load(GOLDEN)
    for i in range(int(g["n_cases"])):
        M, K, N, r, kn, p4, seed, offset = (int(v) for v in g[f"c{i}_meta"])
        p = p4 / 1e4
        t = lambda k: torch.from_numpy(g[f"c{i}_{k}"])
        case = dict(x=t("x"), W=t("W"), b=t("b"), A=t("A"), B=t("B"), dy=t("dy"))
        out, _ = _run_fwd_bwd(pkg, M, K, N, r, p, bool(kn), torch.bfloat16, True, seed=seed, offset=offset, case=case)
        for name, got in (("y", out["Y"]), ("dx", out["dX"]), ("dA", out["dA"]), ("dB", out["dB"])):
            assert relerr(got, t(name)) < TOL and relerr_vec(got, t(name)) < TOL, (i, name)


def test_step0_known_answers(pkg):
    """B = 0 (peft's init): Y equals the base projection bit for bit, dA is exactly zero."""
    M, K, N, r, p = 256, 768, 2304, 8, 0.05
    out, ref = _run_fwd_bwd(pkg, M, K, N, r, p, True, torch.bfloat16, True, b_zero=True)
    base = pkg.binding.linear_base(out["d"]["x"], out["d"]["W"], out["d"]["b"], True)
    assert torch.equal(out["Y"], base)
    assert float(out["dA"].abs().max()) == 0.0
    assert relerr(out["dB"], ref["dB"]) < TOL


def test_merged_weight_equivalence(pkg):
    """p = 0: Y == base(x; W + s B A) up to 16-bit rounding of the merged weight."""
    M, K, N, r = 512, 2560, 2560, 8
    out, _ = _run_fwd_bwd(pkg, M, K, N, r, 0.0, False, torch.bfloat16, True)
    d = out["d"]
    Wm = (d["W"].float() + out["s"] * d["B"] @ d["A"]).to(torch.bfloat16)
    ym = pkg.binding.linear_base(d["x"], Wm, d["b"], False)
    assert relerr(out["Y"], ym.float()) < TOL


@pytest.mark.parametrize("M", [256, 100])  # two row splits folded by the last CTA / one split written by the epilogue
def test_accumulate_flag(pkg, M):
    K, N, r, p = 768, 768, 8, 0.05
    out, ref = _run_fwd_bwd(pkg, M, K, N, r, p, False, torch.bfloat16, True)
    d = out["d"]
    dA = out["dA"].clone()
    dB = out["dB"].clone()
    pkg.binding.linear_bwd(d["dy"], d["x"], d["W"], out["packed"], out["H"], out["keep"], r, out["s"], p, False,
                           out["ws"], need_dx=False, dA=dA, dB=dB, accumulate=True)
    torch.cuda.synchronize()
    assert relerr(dA, 2 * ref["dA"]) < TOL and relerr(dB, 2 * ref["dB"]) < TOL


def test_deterministic_replay(pkg):
    """Same (seed, offset) -> identical bits: what gradient-checkpoint recompute relies on."""
    a, _ = _run_fwd_bwd(pkg, 512, 768, 2304, 8, 0.05, True, torch.bfloat16, True)
    b, _ = _run_fwd_bwd(pkg, 512, 768, 2304, 8, 0.05, True, torch.bfloat16, True)
    for k in ("Y", "H", "dX", "dA", "dB"):
        assert torch.equal(a[k], b[k]), k


def test_rejects_unsupported(pkg):
    x = torch.zeros(16, 12, dtype=torch.bfloat16, device="cuda")
    W = torch.zeros(16, 12, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(RuntimeError, match="multiples of 8"):
        pkg.binding.linear_base(x, W, None, False)
    x = torch.zeros(16, 16, dtype=torch.float32, device="cuda")
    with pytest.raises(RuntimeError, match="bf16 or fp16"):
        pkg.binding.linear_base(x, x, None, False)
    xb = torch.zeros(16, 16, dtype=torch.bfloat16, device="cuda")
    A = torch.zeros(65, 16, device="cuda")
    B = torch.zeros(16, 65, device="cuda")
    ws = torch.empty(1 << 20, dtype=torch.uint8, device="cuda")
    with pytest.raises(RuntimeError, match=r"rank r must be in \[1, 64\]"):
        pkg.binding.linear_fwd(xb, xb, None, A, B, 1.0, 0.0, 0, 0, False, ws)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.binding.linear_base(xb.cpu(), xb.cpu(), None, False)
    small = torch.empty(16, dtype=torch.uint8, device="cuda")
    A8 = torch.zeros(8, 16, device="cuda")
    B8 = torch.zeros(16, 8, device="cuda")
    with pytest.raises(RuntimeError, match="packed-operand block too small"):
        pkg.binding.linear_fwd(xb, xb, None, A8, B8, 1.0, 0.0, 0, 0, False, small)
    _, H, _, packed = pkg.binding.linear_fwd(xb, xb, None, A8, B8, 1.0, 0.0, 0, 0, False)
    with pytest.raises(RuntimeError, match="workspace too small"):
        pkg.binding.linear_bwd(xb, xb, xb, packed, H, None, 8, 1.0, 0.0, False, small)


@pytest.mark.parametrize("M,K,N,w_is_KN", [(16384, 2560, 2560, False), (16384, 768, 2304, True),
                                           (8192, 2560, 10240, False), (8192, 10240, 2560, False)])
def test_full_size_against_fp32_matmul(pkg, M, K, N, w_is_KN):
    """BASELINE-size shapes: too slow for the CPU oracle inside a test, so check against an fp32
    (TF32 off) torch matmul of the same operands on the GPU, and dA/dB by linearity in dY."""
    torch.backends.cuda.matmul.allow_tf32 = False
    dtype, r, p, s = torch.bfloat16, 8, 0.05, 2.0
    c = make_case(M, K, N, r, w_is_KN, dtype)
    d = to_dev(c, dtype)
    ws = torch.empty(pkg.binding.workspace_bytes(M, K, N, r), dtype=torch.uint8, device="cuda")
    Y, H, keep, packed = pkg.binding.linear_fwd(d["x"], d["W"], d["b"], d["A"], d["B"], s, p, 5, 8, w_is_KN)
    dX, dA, dB = pkg.binding.linear_bwd(d["dy"], d["x"], d["W"], packed, H, keep, r, s, p, w_is_KN, ws)
    bits = torch.from_numpy(np.unpackbits(keep.cpu().numpy(), axis=1, bitorder="little")).cuda().float()
    xf, Wf, dyf = d["x"].float(), d["W"].float(), d["dy"].float()
    Wnk = Wf.t() if w_is_KN else Wf  # [N, K]
    xd = xf * bits / (1 - p)
    h = xd @ d["A"].t()
    y_ref = xf @ Wnk.t() + d["b"].float() + s * (h @ d["B"].t())
    dh = s * (dyf @ d["B"])
    dx_ref = dyf @ Wnk + (dh @ d["A"]) * bits / (1 - p)
    for got, want in ((Y, y_ref), (dX, dx_ref), (dA, dh.t() @ xd), (dB, s * (dyf.t() @ h))):
        assert relerr(got, want) < TOL and relerr_vec(got, want) < TOL
    # the keep-bits at full size: spot rows against the numpy generator (the whole mask is 42 M elements)
    rows = [0, 1, 17, M // 2 + 3, M - 1]
    assert np.array_equal(bits[rows].cpu().numpy().astype(bool), philox.keep_rows(rows, K, p, 5, 8))
    # linearity of the adapter gradients in dY: bwd(2 dY) == 2 bwd(dY) up to rounding
    dy2 = (d["dy"].float() * 2).to(dtype)
    _, dA2, dB2 = pkg.binding.linear_bwd(dy2, d["x"], d["W"], packed, H, keep, r, s, p, w_is_KN, ws, need_dx=False)
    assert relerr(dA2, 2 * dA) < 1e-3 and relerr(dB2, 2 * dB) < 1e-3


def test_operands_on_another_device_are_refused(pkg):
    """The library launches on the current device's stream: operands that live elsewhere must be refused at the
    binding (and the autograd nodes switch device themselves, see test_gpu_model.py)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    xb = torch.zeros(16, 16, dtype=torch.bfloat16, device="cuda:1")
    with pytest.raises(RuntimeError, match="current CUDA device"):
        pkg.binding.linear_base(xb, xb, None, False)
    with pytest.raises(RuntimeError, match="on one device"):
        pkg.binding.linear_base(xb, xb.to("cuda:0"), None, False)
    x1 = torch.randn(64, 64, device="cuda:1").to(torch.bfloat16).requires_grad_(True)
    W1 = torch.randn(64, 64, device="cuda:1").to(torch.bfloat16)
    A1, B1 = torch.randn(8, 64, device="cuda:1"), torch.randn(64, 8, device="cuda:1")
    y = pkg.fused_lora_linear(x1, W1, None, A1, B1, 2.0)  # current device is 0: the node switches for the call
    y.sum().backward()
    torch.cuda.synchronize("cuda:1")
    ref = x1.detach().float() @ W1.float().t() + 2.0 * (x1.detach().float() @ A1.t()) @ B1.t()
    assert relerr(y, ref) < TOL and x1.grad is not None
