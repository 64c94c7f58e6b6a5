"""GPU probe: fused cross-entropy forward / backward kernels on the distilgpt2 and OPT logits shapes,
against the HBM floor (forward 2*V bytes per row, backward 4*V)."""
import os, sys, json, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as g

g.build(); pkg = g.load_package(); lib = pkg.binding.lib()
dev = torch.device("cuda:0")
P = lambda t: ctypes.c_void_p(t.data_ptr()); S = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3
for (M, V) in [(16384, 50257), (8192, 50272)]:
    ld = (V + 7) // 8 * 8
    logits = (torch.randn(M, ld, device=dev) * 2).bfloat16()
    labels = torch.randint(0, V, (M,), device=dev)
    lse = torch.empty(M, device=dev); rows = torch.empty(M, device=dev); dl = torch.empty_like(logits)
    scale = torch.full((1,), 1.0 / M, device=dev)
    f = timeit(lambda: lib.lora_ce_fwd(P(logits), ld, P(labels), P(lse), P(rows), M, V, -100, 0, S()))
    b = timeit(lambda: lib.lora_ce_bwd(P(logits), ld, P(labels), P(lse), P(scale), P(dl), ld, M, V, -100, 0, S()))
    t = timeit(lambda: torch.nn.functional.cross_entropy(logits[:, :V].float(), labels))
    gb = M * ld * 2 / 1e9
    print(json.dumps({"M": M, "V": V, "ce_fwd_us": round(f, 1), "fwd_GBps": round(gb / f * 1e6), "ce_bwd_us": round(b, 1),
                      "bwd_GBps": round(2 * gb / b * 1e6), "torch_fp32_ce_fwd_us": round(t, 1)}), flush=True)
