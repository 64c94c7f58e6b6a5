"""GPU probe: every kernel of one LoRA-linear forward + backward timed alone (phase mask), with and without
dropout, at one shape.  Activation buffers rotate through three sets so that no launch finds its input in L2.
LORA_B200_LIB=<experiment build> selects another library (see _build.build_variant); prints one JSON line.

    python scripts/kernel_probe.py [M K N r] [--conv1d]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import __graft_entry__ as g  # noqa: E402

args = [v for v in sys.argv[1:] if not v.startswith("--")]
conv = "--conv1d" in sys.argv
M, K, N, r = (int(v) for v in args[:4]) if len(args) >= 4 else (16384, 2560, 2560, 8)
if not os.environ.get("LORA_B200_LIB"):
    g.build()
pkg = g.load_package()
bind = pkg.binding
dev = torch.device("cuda:0")
s = 16.0 / r
dt = torch.bfloat16
gen = torch.Generator(device="cpu").manual_seed(0)
NSET = 3
xs = [torch.randn(M, K, generator=gen).to(dt).to(dev) for _ in range(NSET)]
dys = [torch.randn(M, N, generator=gen).to(dt).to(dev) for _ in range(NSET)]
W = (torch.randn((K, N) if conv else (N, K), generator=gen) * 0.02).to(dt).to(dev)
b = (torch.randn(N, generator=gen) * 0.02).to(dt).to(dev)
A = ((torch.rand(r, K, generator=gen) * 2 - 1) / K ** 0.5).to(dev)
B = (torch.randn(N, r, generator=gen) * 0.02).to(dev)
ws = torch.zeros(bind.workspace_bytes(M, K, N, r), dtype=torch.uint8, device=dev)
packed = torch.empty(bind.packed_bytes(K, N, r), dtype=torch.uint8, device=dev)
outs = [torch.empty(M, max(K, N), dtype=dt, device=dev) for _ in range(NSET)]
dA, dB = torch.empty(r, K, device=dev), torch.empty(N, r, device=dev)
lib, code = bind.lib(), bind.dtype_code(dt)
st = torch.cuda.current_stream().cuda_stream


def timeit(fn, n=30):
    for i in range(5):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) / n * 1e3, 2)


out = {"lib": os.path.basename(os.environ.get("LORA_B200_LIB", "liblora_b200.so")), "shape": [M, K, N, r], "conv1d": conv}
for p in (0.05, 0.0):
    _, H, keep, _ = bind.linear_fwd(xs[0], W, b, A, B, s, p, 1, 0, conv, packed)
    bind.linear_bwd(dys[0], xs[0], W, packed, H, keep, r, s, p, conv, ws, dA=dA, dB=dB)  # valid dH in the workspace
    kp = keep.data_ptr() if keep is not None else None

    def fwd(i):
        rc = lib.lora_linear_fwd(xs[i % NSET].data_ptr(), W.data_ptr(), b.data_ptr(), A.data_ptr(), B.data_ptr(),
                                 outs[i % NSET].data_ptr(), H.data_ptr(), kp, M, K, N, r, s, p, 1, 0, code, int(conv),
                                 packed.data_ptr(), packed.numel(), st)
        assert rc == 0, bind.last_error()

    def bwd(i):
        rc = lib.lora_linear_bwd(dys[i % NSET].data_ptr(), xs[i % NSET].data_ptr(), W.data_ptr(), packed.data_ptr(),
                                 H.data_ptr(), kp, outs[i % NSET].data_ptr(), dA.data_ptr(), dB.data_ptr(), M, K, N, r, s, p,
                                 code, int(conv), 0, ws.data_ptr(), ws.numel(), st)
        assert rc == 0, bind.last_error()

    tag = f"p{p}"
    if os.environ.get("PROBE_ONLY") == "once":  # one forward + backward per dropout setting (the ncu capture target)
        fwd(1)
        bwd(1)
        torch.cuda.synchronize()
        continue
    if os.environ.get("PROBE_ONLY") == "skinny":  # fast A/B of the HBM-bound kernels
        for name, mask, fn in (("fwd_rowproj", bind.PHASE_ROWPROJ | bind.PHASE_PACK, fwd), ("bwd_rowproj", bind.PHASE_ROWPROJ, bwd),
                               ("colproj", bind.PHASE_COLPROJ, bwd)):
            bind.set_phase_mask(mask)
            try:
                out[f"{name}_us_{tag}"] = timeit(fn)
            finally:
                bind.set_phase_mask(bind.PHASE_ALL)
        continue
    for name, mask, fn in (("fwd_gemm", bind.PHASE_GEMM, fwd), ("bwd_gemm", bind.PHASE_GEMM, bwd),
                           ("fwd_rowproj", bind.PHASE_ROWPROJ | bind.PHASE_PACK, fwd), ("bwd_rowproj", bind.PHASE_ROWPROJ, bwd),
                           ("colproj", bind.PHASE_COLPROJ, bwd)):
        bind.set_phase_mask(mask)
        try:
            out[f"{name}_us_{tag}"] = timeit(fn)
        finally:
            bind.set_phase_mask(bind.PHASE_ALL)
    out[f"fwd_all_us_{tag}"] = timeit(fwd)
    out[f"bwd_all_us_{tag}"] = timeit(bwd)
    out[f"step_us_{tag}"] = timeit(lambda i: (fwd(i), bwd(i)))
print(json.dumps(out))
