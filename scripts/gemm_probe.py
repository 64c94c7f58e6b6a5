"""GPU probe: the tcgen05 GEMM alone (lora_linear_base) vs torch.matmul (cuBLAS, yardstick only)
on the hot-path shapes, with SM clocks sampled under load.  Usage: python scripts/gemm_probe.py"""
import os, sys, json, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as g

g.build()
pkg = g.load_package()
b = pkg.binding
dev = torch.device("cuda:0")


def timeit(fn, n=30, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


import pynvml
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
clk = []
stop = threading.Event()
def sampler():
    while not stop.is_set():
        clk.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0))
        time.sleep(0.005)
threading.Thread(target=sampler, daemon=True).start()

mode = os.environ.get("LORA_B200_GEMM", "2cta")
for (M, K, N, kn) in [(16384, 2560, 2560, False), (16384, 2560, 2560, True), (8192, 2560, 10240, False),
                      (8192, 10240, 2560, False), (16384, 768, 2304, True), (16384, 3072, 768, True),
                      (65536, 2560, 2560, False)]:
    x = torch.randn(M, K, device=dev).bfloat16()
    W = (torch.randn((K, N) if kn else (N, K), device=dev) * 0.02).bfloat16()
    bias = torch.randn(N, device=dev).bfloat16()
    clk.clear()
    t_ours = timeit(lambda: b.linear_base(x, W, bias, kn), n=60)
    c1 = sorted(c for c, _ in clk)[len(clk) // 2] if clk else None
    p1 = max(p for _, p in clk) if clk else None
    clk.clear()
    Wt = W if kn else W.t()
    t_cublas = timeit(lambda: torch.addmm(bias, x, Wt), n=60)
    c2 = sorted(c for c, _ in clk)[len(clk) // 2] if clk else None
    fl = 2.0 * M * K * N
    print(json.dumps({"mode": mode, "M": M, "K": K, "N": N, "w_kn": kn, "ours_ms": round(t_ours, 4),
                      "ours_tflops": round(fl / t_ours / 1e9, 1), "cublas_ms": round(t_cublas, 4),
                      "cublas_tflops": round(fl / t_cublas / 1e9, 1), "sm_mhz_ours": c1, "sm_mhz_cublas": c2,
                      "power_w_max": p1}), flush=True)
stop.set()
