#!/usr/bin/env python
"""Benchmark of the fused LoRA-linear hot path (BASELINE.json configs[1]).

A step = one forward + one backward of ONE LoRA-adapted projection over a batch of M synthetic
tokens (OPT-2.7B q_proj shape by default: 2560 -> 2560, r=8, lora_alpha=16, dropout 0.05, bf16).
`value` is tokens/s with inputs resident in HBM; `e2e` runs the same step through the public
autograd API with pinned HOST buffers (H2D of X and dY, D2H of Y, dX, dA, dB inside the timed
region); `roofline` is the tcgen05 GEMM kernel timed alone, live, with CUDA events;
`cpu_baseline` is the CPU oracle (torch fp32, the reference's engine) on a bounded sample.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    torchrun --nproc-per-node N bench.py --gpus N ...      (one rank per GPU, NCCL)
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import torch

REPO_ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO_ROOT)

METRIC = "lora_linear_fwd_bwd_tokens_per_sec"
UNIT = "tokens/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--tokens", type=int, default=16384, help="M: tokens per step per GPU")
    ap.add_argument("--k", type=int, default=2560)
    ap.add_argument("--n", type=int, default=2560)
    ap.add_argument("--r", type=int, default=8)
    ap.add_argument("--p", type=float, default=0.05)
    ap.add_argument("--dtype", choices=["bf16", "fp16"], default="bf16")
    ap.add_argument("--conv1d", action="store_true", help="GPT-2 Conv1D weight layout [K,N]")
    ap.add_argument("--cpu-tokens", type=int, default=0,
                    help="tokens per step of the CPU legs (0: the reference arm uses --tokens, the cpu_baseline leg of "
                         "the product arm a 4096-token sample)")
    ap.add_argument("--cfg1-ddp", action="store_true",
                    help="with --impl reference under torchrun --nproc-per-node 2: BASELINE configs[0] as a real two-rank "
                         "gloo DDP job on the host cores (every rank works)")
    ap.add_argument("--no-finetune", action="store_true",
                    help="layer workload: skip the short whole-model fine-tune runs that fill the `finetune` key")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sweep", action="store_true", help="run the configs[1] shape sweep, one JSON line per point")
    ap.add_argument("--micro", default="8,32,96",
                    help="opt27b_pp: micro-batches per optimiser step, comma list (96 = deepspeed_pp_zero1_cpu_activ.json:3)")
    ap.add_argument("--workload", choices=["layer", "distilgpt2", "opt27b", "opt27b_pp"], default="layer",
                    help="layer = configs[1] microbench point (default); distilgpt2 / opt27b = whole-model LoRA "
                         "fine-tune step (configs[2] / configs[3]) through get_peft_model + Trainer (+ DDP)")
    ap.add_argument("--batch", type=int, default=0, help="model workloads: sequences per GPU per step")
    ap.add_argument("--seq", type=int, default=0, help="model workloads: sequence length")
    ap.add_argument("--layers", type=int, default=0, help="model workloads: override the number of layers (debug)")
    ap.add_argument("--no-ckpt", action="store_true", help="model workloads: gradient checkpointing off")
    ap.add_argument("--exchange", choices=["p2p", "nccl"], default="p2p",
                    help="N > 1: adapter-gradient bucket averaged by the in-stream peer-memory kernel (default) or by "
                         "NCCL all-reduce (layer workload: on a side stream overlapped with the next step; models: torch DDP)")
    ap.add_argument("--no-graph", action="store_true",
                    help="layer workload: launch every step eagerly instead of replaying one captured CUDA graph per step")
    ap.add_argument("--head", choices=["chunked", "materialised"], default="materialised",
                    help="model workloads: lm_head through the tcgen05 GEMM followed by the fused cross-entropy over "
                         "materialised logits (default, faster), or lm_head + loss in token chunks without [M, V] logits "
                         "(3.3 GB less memory at distilgpt2's 16 x 1024 tokens; its dX GEMM has too few tiles per chunk)")
    ap.add_argument("--head-chunk", type=int, default=0, help="tokens per chunk of the chunked head (0: fit 96 MiB)")
    ap.add_argument("--plain-neighbours", action="store_true",
                    help="model workloads: also move un-adapted frozen projections WITHOUT an activation (OPT q/v: k_proj, "
                         "out_proj, fc2) from cuBLAS onto this library's GEMM (measured 4 % slower on OPT-2.7B)")
    ap.add_argument("--stock-head", action="store_true",
                    help="model workloads: keep torch's lm_head matmul and fp32 cross-entropy (default: lm_head through "
                         "the tcgen05 GEMM + fused cross-entropy, SURVEY 8f rows N2/N3)")
    return ap.parse_args()


def workload_name(a):
    return (f"configs[1] fused LoRA-linear microbench point: fwd+bwd, {a.k}->{a.n} "
            f"({'Conv1D [K,N]' if a.conv1d else 'nn.Linear [N,K]'}), r={a.r}, alpha=16, dropout={a.p}, "
            f"M={a.tokens} tokens/GPU")


def layer_config(a, world):
    """The workload-defining part of `config`, identical in the product arm and the reference arm."""
    M, K, N = a.tokens, a.k, a.n
    return {"workload": workload_name(a), "M": M, "K": K, "N": N, "r": a.r, "lora_alpha": 16, "dropout": a.p,
            "w_layout": "KN" if a.conv1d else "NK", "parallelism": f"dp{world}",
            "l2_policy": "inputs larger than L2 (X, dY, Y, dX, W = %.0f MB per step vs 126 MB L2)" %
                         ((2 * M * K + 2 * M * N + K * N) * 2 / 1e6)}


def flops_fwd_bwd(M, K, N, r):
    return 4.0 * M * K * N + 6.0 * M * r * (K + N)


def flops_gemm_launch(M, K, N, r):
    return 2.0 * M * K * N + 2.0 * M * r * N  # frozen GEMM + the rank-r K-block it carries


def measured_hbm():
    path = os.path.join(REPO_ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


def gpu_unfused_ms(a, x, W, b, A, B, dy, s, p, reps=10):
    """BASELINE.json configs[1] comparator ("vs peft+cuBLAS"; BASELINE.md section 2, B-gpu-unfused): the op sequence
    peft's lora.Linear.forward runs -- base addmm / F.linear, F.dropout, two skinny F.linear, scale, add -- as stock
    torch ops with torch autograd for the backward, same shapes and dtype, on this GPU.  peft itself is not
    installable here.  Two variants: adapters in fp32 (peft's autocast_adapter_dtype default: the input is cast to
    fp32 for the adapter branch) and adapters already in the 16-bit compute dtype (the cheaper configuration)."""
    import torch.nn.functional as F

    out = {}
    for tag, adt in (("fp32_adapters", torch.float32), ("16bit_adapters", x.dtype)):
        Ap = A.to(adt).clone().requires_grad_(True)
        Bp = B.to(adt).clone().requires_grad_(True)
        xr = x.clone().requires_grad_(True)

        def step():
            base = torch.addmm(b, xr, W) if a.conv1d else F.linear(xr, W, b)
            xa = xr.to(adt)
            y = base + (F.linear(F.linear(F.dropout(xa, p, True), Ap), Bp) * s).to(base.dtype)
            y.backward(dy)
            xr.grad = Ap.grad = Bp.grad = None

        for _ in range(3):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            step()
        e1.record()
        torch.cuda.synchronize()
        out[tag + "_ms"] = e0.elapsed_time(e1) / reps
        del Ap, Bp, xr
    torch.cuda.empty_cache()
    return out


def measured_peaks():
    path = os.path.join(REPO_ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return float(d["bf16_tflops"]), float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured"
    return 1590.0, 1400.0, "fallback"


class ClockSampler:
    """Samples SM clock / throttle reasons through NVML while the timed region runs."""

    REASONS = {0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index, enabled=True):
        # NVML start-up is slow and serialises across processes on a multi-GPU box: construct this BEFORE the
        # barrier that precedes a timed region (never between the barrier and the first event), and only on the
        # rank that reports (enabled = rank 0), so that ranks enter the timed loop together.
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        self.nv = None
        if not enabled:
            return
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.004)

    def __enter__(self):
        if self.nv is not None:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t is not None:
            self._t.join()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def make_inputs(a, M, device, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    K, N, r = a.k, a.n, a.r
    x = torch.randn(M, K, generator=g)
    W = torch.randn((K, N) if a.conv1d else (N, K), generator=g) * 0.02
    b = torch.randn(N, generator=g) * 0.02
    A = (torch.rand(r, K, generator=g) * 2 - 1) / (K ** 0.5)
    B = torch.randn(N, r, generator=g) * 0.02  # non-zero, else dA == 0 and the low-rank path idles
    dy = torch.randn(M, N, generator=g)
    cast = lambda t: t.to(dtype).to(device)
    return cast(x), cast(W), cast(b), A.to(device), B.to(device), cast(dy)


# ----------------------------------------------------------------------------------------------
# CPU legs (oracle = torch fp32 CPU restatement of peft's LoRA linear; see oracle/lora_ref.py)
# ----------------------------------------------------------------------------------------------
def cpu_step_time(a, tokens, steps, warmup):
    from oracle import lora_ref, philox

    x, W, b, A, B, dy = make_inputs(a, tokens, "cpu", torch.float32, 1234)
    keep = torch.from_numpy(philox.keep_mask(tokens, a.k, a.p, 1, 0)) if a.p > 0 else None
    s = 16.0 / a.r
    for _ in range(warmup):
        lora_ref.lora_linear_fwd_bwd(x, W, b, A, B, s, dy, keep, a.p, a.conv1d)
    t0 = time.perf_counter()
    for _ in range(steps):
        lora_ref.lora_linear_fwd_bwd(x, W, b, A, B, s, dy, keep, a.p, a.conv1d)
    return (time.perf_counter() - t0) / steps


def use_all_host_threads() -> int:
    """The CPU arms run on every core this process may use.  torchrun exports OMP_NUM_THREADS=1 to its workers,
    which would otherwise pin the reference arm of an N > 1 launch to a single thread."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    torch.set_num_threads(max(1, n))
    return torch.get_num_threads()


def run_reference(a, rank):
    """--impl reference: the reference's own CPU implementation of the path.  peft/accelerate/
    deepspeed cannot be installed here, so this is the oracle port on all host threads, on the SAME
    workload as the product arm (same M, K, N, r, p; one step = forward + backward of the layer over M tokens)."""
    if rank != 0:
        return
    cores = use_all_host_threads()
    tokens = a.cpu_tokens or a.tokens
    t = cpu_step_time(a, tokens, a.steps, max(1, min(a.warmup, 3)))
    v = tokens / t
    sample = f"{tokens} tokens/step x {a.steps} steps of the same layer shape, torch fp32 CPU, {cores} threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": layer_config(a, a.gpus),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ----------------------------------------------------------------------------------------------
# GPU legs
# ----------------------------------------------------------------------------------------------
def exchange_check(pkg, exch, n, device):
    """N > 1, before timing: the peer-memory all-reduce against NCCL on the same random bucket.  Returns the
    dict that goes on the JSON line; raises (rc != 0 for the run) when the exchange is wrong."""
    import torch.distributed as dist

    rank, world = dist.get_rank(), dist.get_world_size()
    g = torch.Generator(device="cpu").manual_seed(4242 + rank)
    worst = 0.0
    same = True
    for size in (n, max(4, n // 3 * 2 + 1)):  # the bucket's own size and a ragged one
        src = torch.randn(size, generator=g).to(device)
        ours = src.clone()
        for mode in (0, 1, 2) if size % 4 == 0 else (0, 1):
            ours.copy_(src)
            exch.allreduce_avg_(ours, mode=mode)
            ref = src.clone()
            dist.all_reduce(ref, op=dist.ReduceOp.SUM)
            ref /= world
            worst = max(worst, float((ours - ref).abs().max()))
            gathered = [torch.empty_like(ours) for _ in range(world)]
            dist.all_gather(gathered, ours)
            same = same and all(torch.equal(gathered[0], t) for t in gathered[1:])
    out = {"max_abs_err_vs_nccl": worst, "bitwise_equal_across_ranks": bool(same), "values": n,
           "modes": "auto, one-shot, two-shot"}
    if not same or not worst < 1e-5:
        raise RuntimeError(f"peer-memory exchange check failed: {out}")
    return out


def run_point(a, pkg, rank, world, device, with_e2e=True, with_kernels=True, with_unfused=True):
    bind = pkg.binding
    dtype = torch.bfloat16 if a.dtype == "bf16" else torch.float16
    M, K, N, r, p = a.tokens, a.k, a.n, a.r, a.p
    s = 16.0 / r
    x, W, b, A, B, dy = make_inputs(a, M, device, dtype, 1234 + rank)
    ws = torch.zeros(bind.workspace_bytes(M, K, N, r), dtype=torch.uint8, device=device)
    packed = torch.empty(bind.packed_bytes(K, N, r), dtype=torch.uint8, device=device)
    Ybuf = torch.empty(M, N, dtype=dtype, device=device)
    nb = r * K + N * r
    # DDP-style flat adapter-gradient buckets (gradient_as_bucket_view): the kernels write dA|dB straight into the
    # bucket.  Two buckets alternate: the all-reduce of step i runs on a side stream under step i+1's forward, the
    # way DDP overlaps a bucket's reduction with the backward that is still running (finetune_lora_distilgpt2.py:142).
    buckets = [torch.zeros(nb, device=device) for _ in range(2)]
    dist_on = world > 1
    exch, xcheck = None, None
    if dist_on:
        import torch.distributed as dist
        if a.exchange == "p2p":
            exch = pkg.comm.try_peer_exchange(nb)  # None (collectively) if CUDA IPC is not available
            if exch is not None:
                xcheck = exchange_check(pkg, exch, nb, device)
    main = torch.cuda.current_stream()
    side = torch.cuda.Stream()
    rng_dev = torch.zeros(1, dtype=torch.int64, device=device)

    def compute(bkt):
        """forward + backward of the layer; gradients land in the bucket."""
        dA, dB = bkt[: r * K].view(r, K), bkt[r * K:].view(N, r)
        _, H, keep, _ = bind.linear_fwd(x, W, b, A, B, s, p, 1, 0, a.conv1d, packed, out=Ybuf, offset_dev=rng_dev)
        bind.linear_bwd(dy, x, W, packed, H, keep, r, s, p, a.conv1d, ws, need_dx=True, dA=dA, dB=dB)
        rng_dev.add_(4)

    def reduce_(bkt):
        if exch is not None:
            exch.allreduce_avg_(bkt)
        else:
            dist.all_reduce(bkt, op=dist.ReduceOp.AVG)

    def step_body(parity):
        """One step: compute into bucket[parity] while the OTHER bucket (previous step's gradients) is averaged on
        the side stream.  Every step issues exactly one exchange; the last one is drained inside the timed region."""
        if dist_on:
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                reduce_(buckets[parity ^ 1])
        compute(buckets[parity])
        if dist_on:
            torch.cuda.current_stream().wait_stream(side)

    def barrier():
        if dist_on:
            dist.barrier()
        torch.cuda.synchronize()

    # One step = one CUDA-graph replay (launch-bound inner loop captured once; 8 processes launching 8 kernels
    # every 0.4 ms from Python otherwise sit on the critical path).  The dropout offset advances on the device
    # inside the graph (lora_linear_fwd_devrng), the exchange kernel keeps its call counter on the device, so
    # every replay draws a fresh mask and is a complete step.  Two graphs, one per bucket parity.  NCCL mode stays
    # eager (its collectives are not captured here).
    graphs = None
    if not a.no_graph and (exch is not None or not dist_on):
        try:
            warm = torch.cuda.Stream()
            warm.wait_stream(main)
            with torch.cuda.stream(warm):
                for i in range(4):
                    step_body(i & 1)
            main.wait_stream(warm)
            torch.cuda.synchronize()
            graphs = []
            for parity in (0, 1):
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    step_body(parity)
                graphs.append(g)
        except Exception as e:  # noqa: BLE001
            print(f"[bench] CUDA-graph capture failed ({e!r}); running eagerly", file=sys.stderr)
            graphs = None
            if dist_on:
                raise  # ranks must not diverge between graph and eager paths

    def run_step(i):
        if graphs is not None:
            graphs[i & 1].replay()
        else:
            step_body(i & 1)

    def drain(i_next):
        """the exchange of the last computed bucket (step i_next - 1), which no later step carried"""
        if dist_on:
            reduce_(buckets[(i_next - 1) & 1])

    for i in range(a.warmup):
        run_step(i)
    drain(a.warmup)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(torch.cuda.current_device(), enabled=rank == 0) as clk:
        barrier()
        e0.record()
        for i in range(a.steps):
            run_step(i)
        drain(a.steps)  # every step's all-reduce is inside the timed region
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    if dist_on:
        t = torch.tensor([ms], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_step = ms / a.steps
    value = world * M / (ms_step * 1e-3)
    graph = graphs

    # ---- every kernel of the step alone, live, CUDA events on this stream ----
    # lora_debug_set_phase_mask picks which kernels a call launches; each is timed over n_rep launches that rotate
    # through three sets of activation buffers (3 x 84 MB per operand > the 126 MB L2), so no launch finds its
    # input resident from the previous one.
    burst, sustained, src = measured_peaks()
    hbm_peak, _ = measured_hbm()
    esz = 2
    rp = bind.rpad(r)
    kernels = []
    roofline = None
    if with_kernels:
        nset = 3
        xs = [x] + [x.clone() for _ in range(nset - 1)]
        dys = [dy] + [dy.clone() for _ in range(nset - 1)]
        outs = [torch.empty(M, max(K, N), dtype=dtype, device=device) for _ in range(nset)]
        _, H, keep, _ = bind.linear_fwd(x, W, b, A, B, s, p, 1, 0, a.conv1d, packed, out=Ybuf)
        dAk, dBk = torch.empty(r, K, device=device), torch.empty(N, r, device=device)
        bind.linear_bwd(dy, x, W, packed, H, keep, r, s, p, a.conv1d, ws, need_dx=True, dA=dAk, dB=dBk)  # valid dH in ws
        lib, code, st = bind.lib(), bind.dtype_code(dtype), torch.cuda.current_stream().cuda_stream
        kp = keep.data_ptr() if keep is not None else None

        def fwd_call(i):
            rc = lib.lora_linear_fwd(xs[i % nset].data_ptr(), W.data_ptr(), b.data_ptr(), A.data_ptr(), B.data_ptr(),
                                     outs[i % nset].data_ptr(), H.data_ptr(), kp, M, K, N, r, s, p, 1, 0, code,
                                     int(a.conv1d), packed.data_ptr(), packed.numel(), st)
            assert rc == 0, bind.last_error()

        def bwd_call(i):
            rc = lib.lora_linear_bwd(dys[i % nset].data_ptr(), xs[i % nset].data_ptr(), W.data_ptr(), packed.data_ptr(),
                                     H.data_ptr(), kp, outs[i % nset].data_ptr(), dAk.data_ptr(), dBk.data_ptr(), M, K, N,
                                     r, s, p, code, int(a.conv1d), 0, ws.data_ptr(), ws.numel(), st)
            assert rc == 0, bind.last_error()

        def time_phase(mask, call):
            n_rep = max(20, min(a.steps, 50))
            bind.set_phase_mask(mask)
            try:
                for i in range(5):
                    call(i)
                torch.cuda.synchronize()
                g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                g0.record()
                for i in range(n_rep):
                    call(i)
                g1.record()
                torch.cuda.synchronize()
                return g0.elapsed_time(g1) / n_rep
            finally:
                bind.set_phase_mask(bind.PHASE_ALL)

        keep_b = M * K // 8 if p > 0 else 0
        gemm_flops = flops_gemm_launch(M, K, N, r)
        spec = [  # name, phase, call, bound, algorithmic flops or bytes per launch
            ("gemm_lora_kernel forward (Y = X W^T + b + H (sB)^T)", bind.PHASE_GEMM, fwd_call, "tensor", gemm_flops),
            ("gemm_lora_kernel backward (dX = dY W + keep/(1-p) .* (dH A)%s)" % (", masked epilogue" if p > 0 else ""),
             bind.PHASE_GEMM, bwd_call, "tensor", 2.0 * M * K * N + 2.0 * M * r * K),
            ("rowproj_kernel forward (H = drop(X) A^T, draws and stores the keep-bits%s)" %
             (", packs the adapters" if rp == 16 else ""), bind.PHASE_ROWPROJ | bind.PHASE_PACK, fwd_call,
             "hbm", M * K * esz + keep_b + M * rp * esz),
            ("rowproj_kernel backward (dH = dY sB)", bind.PHASE_ROWPROJ, bwd_call, "hbm", M * N * esz + M * rp * esz),
            ("colproj_umma_kernel (dB = s dY^T H and dA = dH^T drop(X), one launch, tensor memory)", bind.PHASE_COLPROJ, bwd_call, "hbm",
             M * N * esz + M * K * esz + keep_b + 2 * M * rp * esz),
        ]
        if rp != 16:  # padded rank 16: the forward row projection packs as a side job, there is no pack launch
            spec.append(("pack_adapters_kernel (once per forward)", bind.PHASE_PACK, fwd_call, "launch", 0))
        for name, mask, call, bound, work in spec:
            ms_k = time_phase(mask, call)
            ent = {"kernel": name, "bound": bound, "kernel_ms": ms_k}
            if bound == "tensor":
                ent.update(achieved=work / (ms_k * 1e-3) / 1e12, peak=burst, unit="TFLOP/s")
                ent["frac"] = ent["achieved"] / burst
            elif bound == "hbm":
                ent.update(achieved=work / (ms_k * 1e-3) / 1e9, peak=hbm_peak, unit="GB/s")
                ent["frac"] = ent["achieved"] / hbm_peak
            kernels.append(ent)
        del xs, dys, outs
        traffic = None
        tpath = os.path.join(REPO_ROOT, "profiles", "roofline_traffic.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                traffic = json.load(f).get(f"{M}x{K}x{N}")
        step_tf = flops_fwd_bwd(M, K, N, r) / (ms_step * 1e-3) / 1e12
        top = kernels[0]
        # `roofline` = the dominant kernel (forward GEMM launch) as the contract asks; `kernels` lists every launch
        # of the step, the slower masked backward GEMM and the HBM-bound skinny kernels included, and `step_*` is
        # the whole fwd+bwd step (all seven launches) against the same peaks.
        roofline = {"kernel": "lb::gemm::" + top["kernel"], "bound": "tensor", "achieved": top["achieved"],
                    "peak": burst, "unit": "TFLOP/s", "frac": top["frac"], "traffic": traffic,
                    "peak_source": f"{src} burst bf16 / HBM copy (MEASURED_PEAKS.json)", "kernel_ms": top["kernel_ms"],
                    "kernels": kernels, "kernels_ms_sum": sum(k["kernel_ms"] for k in kernels),
                    "step_tflops": step_tf, "step_frac_of_burst": step_tf / burst,
                    "step_frac_of_sustained": step_tf / sustained}

    # ---- configs[1]'s comparator: the same layer as stock torch ops on cuBLAS (what peft's lora.Linear runs) ----
    unfused = gpu_unfused_ms(a, x, W, b, A, B, dy, s, p) if with_unfused else None

    # ---- e2e: public autograd API, pinned host buffers, copies inside the timed region ----
    e2e = None
    if with_e2e:
        hx = torch.empty(M, K, dtype=dtype).pin_memory().copy_(x.cpu())
        hdy = torch.empty(M, N, dtype=dtype).pin_memory().copy_(dy.cpu())
        hy = torch.empty(M, N, dtype=dtype).pin_memory()
        hdx = torch.empty(M, K, dtype=dtype).pin_memory()
        hda = torch.empty(r, K).pin_memory()
        hdb = torch.empty(N, r).pin_memory()
        Ap, Bp = A.clone().requires_grad_(True), B.clone().requires_grad_(True)

        def e2e_step():
            xd = hx.to(device, non_blocking=True).requires_grad_(True)
            dyd = hdy.to(device, non_blocking=True)
            y = pkg.fused_lora_linear(xd, W, b, Ap, Bp, s, p=p, training=True, w_is_KN=a.conv1d)
            y.backward(dyd)
            if exch is not None:  # the DDP exchange step, on the gradients autograd produced
                exch.allreduce_avg_(Ap.grad)
                exch.allreduce_avg_(Bp.grad)
            hy.copy_(y.detach(), non_blocking=True)
            hdx.copy_(xd.grad, non_blocking=True)
            hda.copy_(Ap.grad, non_blocking=True)
            hdb.copy_(Bp.grad, non_blocking=True)
            Ap.grad = None
            Bp.grad = None

        e2e_steps = max(5, min(a.steps, 20))
        for _ in range(3):
            e2e_step()
        barrier()
        e0.record()
        for _ in range(e2e_steps):
            e2e_step()
        e1.record()
        barrier()
        e2e_ms = e0.elapsed_time(e1) / e2e_steps
        if dist_on:
            t = torch.tensor([e2e_ms], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e2e_ms = float(t.item())
        e2e = {"value": world * M / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": (M * K + M * N) * esz,
               "d2h_bytes_per_step": (M * N + M * K) * esz + (r * K + N * r) * 4, "ms_per_step": e2e_ms,
               "copies": "serial: H2D, compute, D2H on one stream"}

        # The same step with the copies of consecutive steps overlapped, as an input pipeline would run it: step
        # i+1's H2D (copy stream) and step i-1's D2H (second copy stream) run under step i's compute; every step still
        # moves its own inputs in and its own results out inside the timed region (PCIe is full duplex).
        s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
        slots = [{"x": torch.empty(M, K, dtype=dtype, device=device), "dy": torch.empty(M, N, dtype=dtype, device=device),
                  "in": torch.cuda.Event(), "done": torch.cuda.Event(), "out": torch.cuda.Event(), "keep": None}
                 for _ in range(2)]
        hys = [hy, torch.empty_like(hy).pin_memory()]
        hdxs = [hdx, torch.empty_like(hdx).pin_memory()]

        def stage_in(i):
            sl = slots[i & 1]
            with torch.cuda.stream(s_in):
                s_in.wait_event(sl["out"])  # the slot's previous results have left the device
                sl["x"].copy_(hx, non_blocking=True)
                sl["dy"].copy_(hdy, non_blocking=True)
                sl["in"].record(s_in)

        def piped_step(i, prefetch=True):
            sl = slots[i & 1]
            cur = torch.cuda.current_stream()
            cur.wait_event(sl["in"])
            if prefetch:
                stage_in(i + 1)
            xd = sl["x"].detach().requires_grad_(True)
            y = pkg.fused_lora_linear(xd, W, b, Ap, Bp, s, p=p, training=True, w_is_KN=a.conv1d)
            y.backward(sl["dy"])
            if exch is not None:
                exch.allreduce_avg_(Ap.grad)
                exch.allreduce_avg_(Bp.grad)
            ga, gb = Ap.grad, Bp.grad
            Ap.grad = Bp.grad = None
            sl["done"].record(cur)
            sl["keep"] = (y, xd.grad, ga, gb)  # alive until the slot's next use, two steps on
            with torch.cuda.stream(s_out):
                s_out.wait_event(sl["done"])
                hys[i & 1].copy_(y.detach(), non_blocking=True)
                hdxs[i & 1].copy_(xd.grad, non_blocking=True)
                hda.copy_(ga, non_blocking=True)
                hdb.copy_(gb, non_blocking=True)
                sl["out"].record(s_out)

        for sl in slots:
            sl["out"].record()
        stage_in(0)
        for i in range(3):
            piped_step(i)
        torch.cuda.synchronize()
        barrier()
        e0.record()
        for i in range(3, 3 + e2e_steps):
            piped_step(i, prefetch=i + 1 < 3 + e2e_steps)
        torch.cuda.current_stream().wait_stream(s_out)  # the last step's results are on the host
        e1.record()
        barrier()
        piped_ms = e0.elapsed_time(e1) / e2e_steps
        if dist_on:
            t = torch.tensor([piped_ms], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            piped_ms = float(t.item())
        e2e["serial_copies"] = {"value": e2e["value"], "ms_per_step": e2e_ms}
        e2e["pipelined_copies"] = {"value": world * M / (piped_ms * 1e-3), "ms_per_step": piped_ms}
        if piped_ms < e2e_ms:
            e2e.update(value=world * M / (piped_ms * 1e-3), ms_per_step=piped_ms,
                       copies="pipelined: step i+1's H2D and step i-1's D2H on copy streams under step i's compute")
        for sl in slots:
            sl["keep"] = None

        # the same public autograd call with inputs resident in HBM, launched eagerly from Python: what a training
        # script gets per layer call (`value` above is the raw C-ABI sequence replayed as one CUDA graph)
        xres = x.clone().requires_grad_(True)

        def api_step():
            y = pkg.fused_lora_linear(xres, W, b, Ap, Bp, s, p=p, training=True, w_is_KN=a.conv1d)
            y.backward(dy)
            xres.grad = Ap.grad = Bp.grad = None

        for _ in range(3):
            api_step()
        barrier()
        e0.record()
        for _ in range(e2e_steps):
            api_step()
        e1.record()
        barrier()
        api_ms = e0.elapsed_time(e1) / e2e_steps
        e2e["public_api_resident"] = {"value": M / (api_ms * 1e-3), "unit": UNIT + " per GPU", "ms_per_step": api_ms,
                                      "what": "fused_lora_linear(...).backward(dY), eager, inputs in HBM, no exchange"}

    # launches per step: fwd rowproj (+ pack as a side job at padded rank 16, else a pack launch), gemm;
    # bwd rowproj, gemm, colproj (dB+dA, reduction folded in)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": a.dtype, "data": "synthetic",
        "config": layer_config(a, world),
        "run": {"launch": "one CUDA-graph replay per step" if graph is not None else "eager",
                "exchange": ("none" if not dist_on else
                             "peer-memory all-reduce(avg) of the flat adapter-grad bucket every step (csrc/p2p.cuh), on a "
                             "side stream under the next step's forward (two buckets)"
                             if exch is not None else
                             "NCCL all-reduce(avg) of the flat adapter-grad bucket every step, overlapped with the next "
                             "step's compute (two buckets)")},
        "clocks": clk.summary(), "e2e": e2e, "gpu_launches": ((5 if rp == 16 else 6) + (1 if exch is not None else 0)) * a.steps + (1 if exch is not None else 0),
        "roofline": roofline,
    }
    if unfused is not None:
        best = min(unfused.values())
        line["gpu_unfused"] = dict(unfused, what="stock torch ops + autograd on cuBLAS, same layer / shape / dtype "
                                                 "(BASELINE.json configs[1] comparator; peft itself not installable)",
                                   speedup_vs_best_unfused=best / ms_step if world == 1 else None)
    if xcheck is not None:
        line["exchange_check"] = xcheck
    return line


# ----------------------------------------------------------------------------------------------
# whole-model workloads (BASELINE.json configs[2], configs[3]; configs[0] for the CPU arm)
# ----------------------------------------------------------------------------------------------
def model_spec(a):
    """Architecture + LoRA settings of the reference scripts (random init, synthetic tokens)."""
    if a.workload == "distilgpt2":
        from transformers import GPT2Config, GPT2LMHeadModel

        cfg = GPT2Config(n_layer=a.layers or 6, n_embd=768, n_head=12, n_positions=1024, vocab_size=50257)
        targets = ["c_attn", "c_proj"]  # finetune_lora_distilgpt2.py:100
        return GPT2LMHeadModel, cfg, targets, a.batch or 16, a.seq or 1024, 50256
    from transformers import OPTConfig, OPTForCausalLM

    cfg = OPTConfig(hidden_size=2560, ffn_dim=10240, num_hidden_layers=a.layers or 32, num_attention_heads=32,
                    word_embed_proj_dim=2560, vocab_size=50272, max_position_embeddings=2048,
                    do_layer_norm_before=True, activation_function="relu", dropout=0.1)
    targets = ["q_proj", "v_proj"]  # BASELINE.json configs[3]
    return OPTForCausalLM, cfg, targets, a.batch or 4, a.seq or 2048, 1


def model_workload_name(a, B, S, world):
    return (f"{a.workload} architecture (random init) LoRA r=8 alpha=16 dropout=0.05 fine-tune step, "
            f"gradient checkpointing {'off' if a.no_ckpt else 'on'}, batch {B} x seq {S} per GPU, "
            f"{('DDP, adapter bucket averaged by ' + ('the in-stream peer-memory kernel, ' if a.exchange == 'p2p' else 'NCCL all-reduce, ')) if world > 1 else ''}AdamW lr 5e-5 + clip 1.0, "
            f"{'stock torch lm_head + fp32 CE' if a.stock_head else ('chunked lm_head + CE (no [M,V] logits)' if a.head == 'chunked' else 'lm_head on the tcgen05 GEMM + fused CE over materialised logits')}"
            f"{', every frozen projection of the blocks on the tcgen05 GEMM' if a.plain_neighbours and not a.stock_head else ''}")


def run_model(a, pkg, rank, world, device, with_e2e=True):
    import gc

    import torch.distributed as dist

    cls, cfg, targets, B, S, pad_id = model_spec(a)
    dtype = torch.bfloat16 if a.dtype == "bf16" else torch.float16
    torch.manual_seed(0)  # finetune_lora_distilgpt2.py:30
    with torch.device(device):
        model = cls(cfg).to(dtype)
    model.config.use_cache = False
    if not a.no_ckpt:
        model.gradient_checkpointing_enable()  # finetune_lora_distilgpt2.py:93-96
    pm = pkg.get_peft_model(model, pkg.LoraConfig(r=8, lora_alpha=16, target_modules=targets, lora_dropout=0.05,
                                                  bias="none", task_type=pkg.TaskType.CAUSAL_LM))
    n_lora = len(pm.base_model.lora_layers())
    loss_fn = None
    if not a.stock_head:
        # lm_head fused with the loss in token chunks (no [M, V] logits), frozen neighbours of the
        # adapters (c_fc + gelu_new, fc1 + relu, and for OPT q/v the un-targeted k_proj / out_proj / fc2) on this
        # library's tcgen05 GEMM instead of cuBLAS
        if a.head == "chunked":
            loss_fn = pkg.frozen.fuse_lm_head_with_loss(pm, chunk_tokens=a.head_chunk)
        else:
            pkg.frozen.accelerate_lm_head(pm)
            loss_fn = pkg.frozen.fused_causal_lm_loss
        pkg.frozen.accelerate_frozen_neighbours(pm, plain=a.plain_neighbours)
    net = pkg.train.wrap_ddp(pm, device, exchange=a.exchange) if world > 1 else pm
    tr = pkg.train.Trainer(net, pkg.train.TrainConfig(lr=5e-5, accum=1, total_steps=10 ** 6), loss_fn=loss_fn)
    tr._ensure_sched(10 ** 6)
    net.train()
    n_batches = 4
    host = pkg.train.synthetic_batches(n_batches, B, S, cfg.vocab_size, pad_id, seed=1234 + rank)
    host = [{k: v.pin_memory() for k, v in b.items()} for b in host]
    dev_batches = [{k: v.to(device) for k, v in b.items()} for b in host]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident(i):
        loss = tr.micro_step(dev_batches[i % n_batches], sync=True)
        tr.optimizer_step()
        return loss

    def step_e2e(i):
        b = {k: v.to(device, non_blocking=True) for k, v in host[i % n_batches].items()}
        loss = tr.micro_step(b, sync=True)
        tr.optimizer_step()
        return float(loss)  # device -> host read of the step's loss

    def timed(fn, steps, warmup):
        for i in range(warmup):
            fn(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with ClockSampler(torch.cuda.current_device(), enabled=rank == 0) as clk:
            barrier()
            e0.record()
            for i in range(steps):
                fn(i)
            e1.record()
            barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms / steps, clk

    ms_step, clk = timed(step_resident, a.steps, a.warmup)
    e2e_ms = timed(step_e2e, max(3, min(a.steps, 10)), 2)[0] if with_e2e else float("nan")
    tokens = world * B * S
    fwd_calls = n_lora * (1 if a.no_ckpt else 2)  # checkpointing re-runs every block's forward
    line = {
        "metric": "lora_finetune_tokens_per_sec", "value": tokens / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": a.dtype, "data": "synthetic",
        "config": {"workload": model_workload_name(a, B, S, world), "global_batch": world * B, "seq_len": S,
                   "lora_modules": n_lora, "trainable_params": pm.get_nb_trainable_parameters()[0],
                   "parallelism": f"dp{world}", "l2_policy": "activations + weights far larger than L2"},
        "clocks": clk.summary(),
        "e2e": {"value": tokens / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 3 * B * S * 8,
                "d2h_bytes_per_step": 4, "ms_per_step": e2e_ms},
        "gpu_launches": a.steps * (fwd_calls * 3 + n_lora * 3),
    }
    if not with_e2e:
        line["e2e"] = None
    close = getattr(pkg.train, "close_exchange", None)
    if close is not None:
        close(net)
    del tr, net, pm, model, dev_batches
    gc.collect()
    torch.cuda.empty_cache()
    return line


# ----------------------------------------------------------------------------------------------
# pipeline workload (BASELINE.json configs[4]): OPT-2.7B LoRA fp16, 32 layers over WORLD_SIZE stages
# ----------------------------------------------------------------------------------------------
def run_pipeline(a, pkg, rank, world, device):
    """finetune_lora_opt_pp.py:156 (split), :121-153 (embeddings first, head + loss last), :169-175 / :213-224
    (engine + train_batch), deepspeed_pp_zero1_cpu_activ.json:2-3 (micro-batch 1 x 96 accumulation steps): one
    optimiser step = `m` micro-batches of 1 x 2048 tokens through the stages, 1F1B, NCCL send/recv.  Runs every
    m of --micro (comma list) and both exchange modes, reports tokens/s, the measured bubble against
    (p-1)/(m+p-1), bytes per stage boundary and the NVLink rate of the boundary transfer alone."""
    import torch.distributed as dist
    from transformers import OPTConfig, OPTForCausalLM

    if world < 2:
        raise SystemExit("--workload opt27b_pp needs >= 2 ranks (torchrun --nproc-per-node S)")
    dtype = torch.float16  # BASELINE.json configs[4]
    layers, seq, mb = a.layers or 32, a.seq or 2048, a.batch or 1
    cfg = OPTConfig(hidden_size=2560, ffn_dim=10240, num_hidden_layers=layers, num_attention_heads=32,
                    word_embed_proj_dim=2560, vocab_size=50272, max_position_embeddings=2048, do_layer_norm_before=True,
                    activation_function="relu", dropout=0.1)
    torch.manual_seed(0)
    with torch.device("meta"):  # only this stage's share is ever materialised
        full = OPTForCausalLM(cfg).to(dtype)
    full.config.use_cache = False
    targets = ["q_proj", "k_proj", "v_proj", "out_proj", "fc1", "fc2"]  # finetune_lora_opt_pp.py:109
    pm = pkg.get_peft_model(full, pkg.LoraConfig(r=8, lora_alpha=16, lora_dropout=0.05, bias="none",
                                                 target_modules=targets, task_type=pkg.TaskType.CAUSAL_LM))
    loss_fn = None
    if not a.stock_head:
        if a.head == "chunked":
            loss_fn = pkg.frozen.fuse_lm_head_with_loss(pm, chunk_tokens=a.head_chunk)
        else:
            pkg.frozen.accelerate_lm_head(pm)
            loss_fn = pkg.frozen.fused_causal_lm_loss
    stage = pkg.pipeline.OPTStage(pm.base_model.model, rank, world, loss_fn=loss_fn)
    stage.to_empty(device=device)
    gen = torch.Generator(device=device).manual_seed(1 + rank)
    with torch.no_grad():
        for n, prm in stage.named_parameters():
            if "lora_B" in n:
                prm.normal_(0.0, 0.02, generator=gen)  # non-zero, else dA == 0 and the low-rank path idles
            elif prm.dim() > 1:
                prm.normal_(0.0, 0.02, generator=gen)
            elif "layer_norm" in n and n.endswith("weight"):
                prm.fill_(1.0)
            else:
                prm.zero_()
    for n, prm in stage.named_parameters():
        prm.requires_grad_("lora_" in n)

    def barrier():
        dist.barrier()
        torch.cuda.synchronize()

    # the boundary transfer alone: rank 0 -> rank 1, [mb, seq, hidden] fp16
    hid = torch.empty(mb, seq, cfg.hidden_size, dtype=dtype, device=device)
    link_ms = None
    if rank < 2:
        for _ in range(3):
            dist.send(hid, 1) if rank == 0 else dist.recv(hid, 0)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            dist.send(hid, 1) if rank == 0 else dist.recv(hid, 0)
        e1.record()
        torch.cuda.synchronize()
        link_ms = e0.elapsed_time(e1) / 20
    boundary_bytes = hid.numel() * hid.element_size()

    ms_list = [int(v) for v in str(a.micro).split(",")]
    results = []
    clk_summary = None
    for overlap in (True, False):
        eng = pkg.pipeline.PipelineEngine(stage, lr=5e-5, max_grad_norm=1.0, act_dtype=dtype, device=device,
                                          overlap=overlap)
        for m in ms_list:
            batches = pkg.train.synthetic_batches(m, mb, seq, cfg.vocab_size, 1, seed=1234, device=device)

            def step():
                eng.zero_grad()
                return eng.train_batch(batches)

            steps = max(2, min(a.steps, 5))
            for _ in range(max(1, min(a.warmup, 2))):
                step()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with ClockSampler(torch.cuda.current_device(), enabled=rank == 0) as clk:
                barrier()
                e0.record()
                for _ in range(steps):
                    loss = step()
                e1.record()
                barrier()
            t = torch.tensor([e0.elapsed_time(e1) / steps], device=device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            l = torch.tensor([float(loss) if loss is not None else 0.0], device=device)
            dist.all_reduce(l)
            results.append({"m": m, "overlap": overlap, "ms_per_step": float(t), "tokens_per_s": m * mb * seq / (float(t) * 1e-3),
                            "loss": float(l), "steps": steps, "bubble_ideal": (world - 1) / (m + world - 1),
                            "p2p_bytes_this_stage_per_step": eng.p2p_bytes})
            clk_summary = clk.summary()
    # measured bubble: the per-micro-batch time of the slowest stage is the slope of T(m); bubble = 1 - m * slope / T(m)
    for overlap in (True, False):
        rs = sorted((r for r in results if r["overlap"] == overlap), key=lambda r: r["m"])
        if len(rs) >= 2:
            slope = (rs[-1]["ms_per_step"] - rs[-2]["ms_per_step"]) / (rs[-1]["m"] - rs[-2]["m"])
            for r in rs:
                r["ms_per_micro_batch_slope"] = slope
                r["bubble_measured"] = 1.0 - r["m"] * slope / r["ms_per_step"]
    best = max((r for r in results if r["m"] == ms_list[-1]), key=lambda r: r["tokens_per_s"])
    link = torch.tensor([link_ms or 0.0], device=device)
    dist.all_reduce(link, op=dist.ReduceOp.MAX)
    line = {
        "metric": "lora_finetune_tokens_per_sec", "value": best["tokens_per_s"], "unit": UNIT, "n_gpus": world,
        "steps": best["steps"], "warmup": max(1, min(a.warmup, 2)), "ms_per_step": best["ms_per_step"],
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "fp16", "data": "synthetic",
        "config": {"workload": f"OPT-2.7B architecture (random init) LoRA r=8 alpha=16 dropout=0.05 on six targets, "
                               f"pipeline-parallel: {layers} layers over {world} stages, 1F1B, {best['m']} micro-batches of "
                               f"{mb} x {seq} tokens per optimiser step, AdamW + clip 1.0, "
                               f"{'stock torch lm_head + fp32 CE' if a.stock_head else ('chunked lm_head + CE' if a.head == 'chunked' else 'lm_head on the tcgen05 GEMM + fused CE')}",
                   "split": pkg.pipeline.split_layers(layers, world), "parallelism": f"pp{world}", "global_batch": best["m"] * mb,
                   "seq_len": seq, "l2_policy": "activations + weights far larger than L2"},
        "clocks": clk_summary,
        "pipeline": {"runs": results, "boundary_bytes": boundary_bytes, "boundary_transfer_ms": float(link),
                     "boundary_GBps": boundary_bytes / (float(link) * 1e-3) / 1e9 if float(link) > 0 else None,
                     "nvlink_peer_copy_reference_GBps": 770.0},
        "e2e": None, "gpu_launches": None,
    }
    return line


def run_reference_cfg1(a):
    """--impl reference --workload distilgpt2 --cfg1-ddp under `torchrun --nproc-per-node 2`: BASELINE.json
    configs[0] / BASELINE.md section 2 "B-cfg1" exactly -- distilgpt2 architecture (random init), LoRA r=8 alpha=16
    dropout 0.05 on c_attn / c_proj (18 modules, 405 504 trainable parameters), fp32, DDP over gloo with world_size 2,
    seq 128, batch 4 per rank, AdamW lr 5e-5, clip 1.0 (finetune_lora_distilgpt2.py:98-104, 130-145) on the oracle's
    layer.  3 warm-up + 20 timed optimiser steps, then the 50-step loss curve.  EVERY rank works here (the run is the
    two-rank DDP job itself); rank 0 prints the line."""
    import torch.distributed as dist
    from transformers import GPT2Config, GPT2LMHeadModel

    import __graft_entry__ as g
    from oracle import lora_ref

    pkg = g.load_package()
    rank, world, _ = pkg.train.init_distributed("gloo")
    threads = int(os.environ.get("OMP_NUM_THREADS", "0")) or max(1, len(os.sched_getaffinity(0)) // max(world, 1))
    torch.set_num_threads(threads)
    B, S, warm, timed, curve = 4, 128, 3, 20, 50
    torch.manual_seed(0)  # finetune_lora_distilgpt2.py:30
    model = GPT2LMHeadModel(GPT2Config(n_layer=6, n_embd=768, n_head=12, n_positions=1024, vocab_size=50257))
    model.config.use_cache = False
    pm = lora_ref.ref_get_peft_model(model, lora_ref.RefLoraConfig(r=8, lora_alpha=16, lora_dropout=0.05,
                                                                  target_modules=["c_attn", "c_proj"]))
    n_train = sum(p.numel() for p in pm.parameters() if p.requires_grad)
    net = pkg.train.wrap_ddp(pm) if world > 1 else pm
    batches = pkg.train.synthetic_batches(warm + timed + curve, B, S, 50257, 50256, seed=1234 + rank)
    tr = pkg.train.Trainer(net, pkg.train.TrainConfig(lr=5e-5, accum=1, total_steps=warm + timed + curve))
    tr._ensure_sched(warm + timed + curve)
    net.train()

    def step(b):
        loss = tr.micro_step(b, sync=True)
        tr.optimizer_step()
        return float(loss)

    for b in batches[:warm]:
        step(b)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for b in batches[warm:warm + timed]:
        step(b)
    if world > 1:
        dist.barrier()
    wall = time.perf_counter() - t0
    losses = [step(b) for b in batches[warm + timed:]]
    if rank == 0:
        v = world * B * S * timed / wall
        sample = f"{timed} timed optimiser steps after {warm} warm-up, {threads} threads per rank"
        print(json.dumps({
            "impl": "reference", "metric": "lora_finetune_tokens_per_sec", "value": v, "unit": UNIT, "n_gpus": 0,
            "ms_per_step": wall / timed * 1e3, "steps": timed, "warmup": warm, "dtype": "f32", "data": "synthetic",
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "config": {"workload": "BASELINE.json configs[0]: distilgpt2 architecture LoRA (r=8, c_attn/c_proj) DDP on "
                                   f"CPU/gloo world_size={world}, seq {S}, batch {B} per rank, fp32, AdamW lr 5e-5, clip 1.0",
                       "trainable_params": n_train, "parallelism": f"dp{world} (gloo)"},
            "cpu_baseline": {"value": v, "unit": UNIT, "kind": "port", "cores": threads * world, "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
            "loss_curve_50_steps": [round(x, 5) for x in losses],
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_reference_model(a, rank):
    """CPU arm for the model workloads: BASELINE.json configs[0] -- distilgpt2-architecture LoRA (r=8, c_attn/c_proj)
    fp32 on the host cores, synthetic tokens seq 128 batch 4 (the reference's own CPU-runnable case); for opt27b a
    bounded 2-layer slice of the same architecture."""
    if rank != 0:
        return
    use_all_host_threads()
    from oracle import lora_ref

    a2 = argparse.Namespace(**vars(a))
    if a.workload == "opt27b":
        a2.layers = a.layers or 2
    cls, cfg, targets, _, _, pad_id = model_spec(a2)
    B, S = 4, 128
    torch.manual_seed(0)
    model = cls(cfg)
    model.config.use_cache = False
    pm = lora_ref.ref_get_peft_model(model, lora_ref.RefLoraConfig(r=8, lora_alpha=16, target_modules=targets,
                                                                  lora_dropout=0.05))
    import __graft_entry__ as g

    pkg = g.load_package()
    batches = pkg.train.synthetic_batches(max(2, a.steps + 1), B, S, cfg.vocab_size, pad_id, seed=1234)
    lora_ref.train_steps(pm, batches[:1], lr=5e-5)
    t0 = time.perf_counter()
    lora_ref.train_steps(pm, batches[1:1 + a.steps], lr=5e-5)
    t = (time.perf_counter() - t0) / a.steps
    v = B * S / t
    sample = f"batch {B} x seq {S}, {cfg.num_hidden_layers if hasattr(cfg, 'num_hidden_layers') else cfg.n_layer} layers, fp32, {a.steps} steps"
    print(json.dumps({
        "impl": "reference", "metric": "lora_finetune_tokens_per_sec", "value": v, "unit": UNIT, "n_gpus": a.gpus,
        "steps": a.steps, "warmup": a.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{a.workload} architecture LoRA fine-tune step on CPU (configs[0] style)", "sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0,
    }))


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        if a.cfg1_ddp:
            run_reference_cfg1(a)
            return
        if a.workload == "layer":
            run_reference(a, rank)
        else:
            run_reference_model(a, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the fused LoRA kernels have no CPU fallback "
                         "(use --impl reference for the CPU reference arm)")
    import __graft_entry__ as g

    g.build()
    pkg = g.load_package()
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        import datetime

        dist.init_process_group("nccl", device_id=device,
                                timeout=datetime.timedelta(seconds=int(os.environ.get("LORA_B200_DIST_TIMEOUT_S", "600"))))
    try:
        if a.workload == "opt27b_pp":
            line = run_pipeline(a, pkg, rank, world, device)
            if rank == 0:
                print(json.dumps(line), flush=True)
            return
        if a.workload != "layer":
            line = run_model(a, pkg, rank, world, device)
            if rank == 0:
                print(json.dumps(line), flush=True)
            return
        if a.sweep:
            # BASELINE.json configs[1]: d_model 768 / 2560 shapes x r 8/16/64 x tokens 4K-64K, fwd+bwd, each point
            # with the stock-torch (cuBLAS) comparator beside it
            shapes = [(768, 2304, True), (768, 768, True), (3072, 768, True), (2560, 2560, False),
                      (2560, 10240, False), (10240, 2560, False)]
            for (K, N, conv) in shapes:
                for r in (8, 16, 64):
                    for M in (4096, 8192, 16384, 32768, 65536):
                        a.k, a.n, a.conv1d, a.r, a.tokens = K, N, conv, r, M
                        line = run_point(a, pkg, rank, world, device, with_e2e=False)
                        if rank == 0:
                            rf = line["roofline"]
                            rf["kernels"] = [{"kernel": k["kernel"].split(" ")[0] + (" fwd" if "forward" in k["kernel"] else
                                                                                   " bwd" if "backward" in k["kernel"] else ""),
                                              "kernel_ms": round(k["kernel_ms"], 5), "frac": round(k.get("frac", 0.0), 4)}
                                             for k in rf["kernels"]]
                            print(json.dumps({k: line[k] for k in ("value", "ms_per_step", "config", "roofline", "gpu_unfused")}),
                                  flush=True)
                        torch.cuda.empty_cache()
            return
        line = run_point(a, pkg, rank, world, device)
        if not a.no_finetune:
            # The first half of BASELINE.json's metric -- "LoRA fine-tune tokens/sec at 1/2/4/8 B200" -- on the same
            # line: a short run of each whole-model workload (configs[2], configs[3]) at this N.
            line["finetune"] = {}
            for wl in ("distilgpt2", "opt27b"):
                a2 = argparse.Namespace(**vars(a))
                a2.workload, a2.steps, a2.warmup = wl, 8, 3
                torch.cuda.empty_cache()
                m = run_model(a2, pkg, rank, world, device, with_e2e=False)
                line["finetune"][wl] = {k: m[k] for k in ("metric", "value", "unit", "ms_per_step", "steps", "warmup", "clocks")}
                line["finetune"][wl]["config"] = m["config"]
        if rank == 0:
            if world == 1 and not a.no_cpu_baseline:
                cores = use_all_host_threads()
                tokens = a.cpu_tokens or 4096
                t = cpu_step_time(a, tokens, 5, 1)
                line["cpu_baseline"] = {"value": tokens / t, "unit": UNIT, "cores": cores, "kind": "port",
                                        "sample": f"{tokens} tokens/step x 5 steps, same layer shape, "
                                                  "oracle/lora_ref.py (torch fp32 CPU)"}
            print(json.dumps(line), flush=True)
    finally:
        if world > 1:
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
