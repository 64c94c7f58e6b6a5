"""GPU probe of epilogue-bound (short-K) shapes: plain GEMM with / without bias, activation epilogue
with / without the pre-activation store, against torch addmm (yardstick only)."""
import os, sys, json, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as g

g.build()
pkg = g.load_package()
b = pkg.binding
dev = torch.device("cuda:0")
P = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())
S = lambda: ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def timeit(fn, n=50, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


lib = b.lib()
for (M, K, N, kn) in [(16384, 768, 2304, True), (16384, 768, 3072, True), (16384, 768, 768, True), (16384, 3072, 768, True),
                      (16384, 2304, 768, False), (8192, 2560, 10240, False), (16384, 2560, 2560, False)]:
    x = torch.randn(M, K, device=dev).bfloat16()
    W = (torch.randn((K, N) if kn else (N, K), device=dev) * 0.02).bfloat16()
    bias = torch.randn(N, device=dev).bfloat16()
    y = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    z = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    r = {"M": M, "K": K, "N": N}
    r["plain_bias_us"] = timeit(lambda: lib.lora_frozen_linear_fwd(P(x), K, P(W), P(bias), P(y), N, M, K, N, 0, int(kn), S()))
    r["plain_nobias_us"] = timeit(lambda: lib.lora_frozen_linear_fwd(P(x), K, P(W), None, P(y), N, M, K, N, 0, int(kn), S()))
    r["gelu_z_us"] = timeit(lambda: lib.lora_frozen_linear_act_fwd(P(x), K, P(W), P(bias), P(y), N, P(z), N, M, K, N, 2, 0, int(kn), S()))
    r["gelu_noz_us"] = timeit(lambda: lib.lora_frozen_linear_act_fwd(P(x), K, P(W), P(bias), P(y), N, None, N, M, K, N, 2, 0, int(kn), S()))
    r["relu_noz_us"] = timeit(lambda: lib.lora_frozen_linear_act_fwd(P(x), K, P(W), P(bias), P(y), N, None, N, M, K, N, 1, 0, int(kn), S()))
    Wt = W if kn else W.t()
    r["addmm_us"] = timeit(lambda: torch.addmm(bias, x, Wt))
    r["mm_us"] = timeit(lambda: torch.mm(x, Wt))
    r = {k: (round(v, 1) if isinstance(v, float) else v) for k, v in r.items()}
    r["ideal_us_1440TF"] = round(2.0 * M * K * N / 1440e12 * 1e6, 1)
    print(json.dumps(r), flush=True)
