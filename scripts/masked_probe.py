"""GPU probe: the backward dX GEMM alone (phase mask = GEMM only) with dropout (masked epilogue) and without
(LoRA K-block), headline shape; LORA_B200_DEBUG_MASKED_R in the environment cuts the rank-r loop (timing only)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as g
g.build(); pkg = g.load_package(); bind = pkg.binding
dev = torch.device("cuda:0")
M, K, N, r, s = 16384, 2560, 2560, 8, 2.0
x = torch.randn(M, K, device=dev).bfloat16(); W = (torch.randn(N, K, device=dev) * 0.02).bfloat16()
b = torch.randn(N, device=dev).bfloat16(); A = torch.randn(r, K, device=dev) * 0.02; B = torch.randn(N, r, device=dev) * 0.02
dy = torch.randn(M, N, device=dev).bfloat16()
ws = torch.empty(bind.workspace_bytes(M, K, N, r), dtype=torch.uint8, device=dev)
def timeit(fn, n=40):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
out = {"debug_masked_r": os.environ.get("LORA_B200_DEBUG_MASKED_R")}
for p in (0.05, 0.0):
    Y, H, keep, xd = bind.linear_fwd(x, W, b, A, B, s, p, 1, 0, False, ws, save_xd=True)
    xin = xd if xd is not None else x
    bind.linear_bwd(dy, xin, W, A, B, H, keep, s, p, False, ws, x_is_dropped=xd is not None)  # valid packed operands + dH
    bind.set_phase_mask(bind.PHASE_GEMM)
    try:
        out[f"bwd_gemm_us_p{p}"] = round(timeit(lambda: bind.linear_bwd(dy, xin, W, A, B, H, keep, s, p, False, ws, x_is_dropped=xd is not None)), 1)
        out[f"fwd_gemm_us_p{p}"] = round(timeit(lambda: bind.linear_fwd(x, W, b, A, B, s, p, 1, 0, False, ws)), 1)
    finally:
        bind.set_phase_mask(bind.PHASE_ALL)
print(json.dumps(out))
