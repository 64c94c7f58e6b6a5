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
    ap.add_argument("--cpu-tokens", type=int, default=2048, help="tokens per step of the CPU legs (bounded sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sweep", action="store_true", help="run the configs[1] shape sweep, one JSON line per point")
    ap.add_argument("--workload", choices=["layer", "distilgpt2", "opt27b"], default="layer",
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
    ap.add_argument("--stock-head", action="store_true",
                    help="model workloads: keep torch's lm_head matmul and fp32 cross-entropy (default: lm_head through "
                         "the tcgen05 GEMM + fused cross-entropy, SURVEY 8f rows N2/N3)")
    return ap.parse_args()


def workload_name(a):
    return (f"configs[1] fused LoRA-linear microbench point: fwd+bwd, {a.k}->{a.n} "
            f"({'Conv1D [K,N]' if a.conv1d else 'nn.Linear [N,K]'}), r={a.r}, alpha=16, dropout={a.p}, "
            f"M={a.tokens} tokens/GPU")


def flops_fwd_bwd(M, K, N, r):
    return 4.0 * M * K * N + 6.0 * M * r * (K + N)


def flops_gemm_launch(M, K, N, r):
    return 2.0 * M * K * N + 2.0 * M * r * N  # frozen GEMM + the rank-r K-block it carries


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
    deepspeed cannot be installed here, so this is the oracle port on all host threads."""
    if rank != 0:
        return
    cores = use_all_host_threads()
    t = cpu_step_time(a, a.cpu_tokens, a.steps, max(1, min(a.warmup, 3)))
    v = a.cpu_tokens / t
    sample = f"{a.cpu_tokens} tokens/step x {a.steps} steps of the same layer shape, torch fp32 CPU"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": {"workload": workload_name(a), "sample": sample},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ----------------------------------------------------------------------------------------------
# GPU legs
# ----------------------------------------------------------------------------------------------
def run_point(a, pkg, rank, world, device, with_e2e=True):
    bind = pkg.binding
    dtype = torch.bfloat16 if a.dtype == "bf16" else torch.float16
    M, K, N, r, p = a.tokens, a.k, a.n, a.r, a.p
    s = 16.0 / r
    x, W, b, A, B, dy = make_inputs(a, M, device, dtype, 1234 + rank)
    ws = torch.empty(bind.workspace_bytes(M, K, N, r), dtype=torch.uint8, device=device)
    # DDP-style flat adapter-gradient buckets (gradient_as_bucket_view): the kernels write dA|dB straight
    # into the bucket; two buckets alternate so that the all-reduce of step i overlaps step i+1's compute,
    # the way DDP overlaps the bucket reduce with the remaining backward.
    buckets = [torch.empty(r * K + N * r, device=device) for _ in range(2)]
    pending = [None, None]
    dist_on = world > 1
    exch = None
    if dist_on:
        import torch.distributed as dist
        if a.exchange == "p2p":
            exch = pkg.comm.try_peer_exchange(r * K + N * r)  # None (collectively) if CUDA IPC is not available

    def step(i):
        bkt = buckets[i & 1]
        if pending[i & 1] is not None:
            pending[i & 1].wait()  # stream-level wait: the bucket's previous all-reduce has finished
            pending[i & 1] = None
        dA, dB = bkt[: r * K].view(r, K), bkt[r * K:].view(N, r)
        Y, H, keep, xd = bind.linear_fwd(x, W, b, A, B, s, p, 1, 4 * i, a.conv1d, ws, save_xd=True)
        bind.linear_bwd(dy, xd if xd is not None else x, W, A, B, H, keep, s, p, a.conv1d, ws, need_dx=True, dA=dA,
                        dB=dB, x_is_dropped=xd is not None)
        if exch is not None:  # the path's one exchange step under DDP: average the adapter gradients, in stream
            exch.allreduce_avg_(bkt)
        elif dist_on:
            pending[i & 1] = dist.all_reduce(bkt, op=dist.ReduceOp.AVG, async_op=True)
        return Y

    def drain():
        for j in range(2):
            if pending[j] is not None:
                pending[j].wait()
                pending[j] = None

    def barrier():
        if dist_on:
            dist.barrier()
        torch.cuda.synchronize()

    # One step = one CUDA-graph replay (launch-bound inner loop captured once; 8 processes launching 8 kernels
    # every 0.4 ms from Python otherwise sit on the critical path).  The dropout offset advances on the device
    # inside the graph (lora_linear_fwd_devrng), the exchange kernel keeps its call counter on the device, so
    # every replay draws a fresh mask and is a complete step.  NCCL mode stays eager (side-stream overlap).
    graph = None
    rng_dev = torch.zeros(1, dtype=torch.int64, device=device)
    bkt0 = buckets[0]
    dA0, dB0 = bkt0[: r * K].view(r, K), bkt0[r * K:].view(N, r)

    def graph_step():
        Y, H, keep, xd = bind.linear_fwd(x, W, b, A, B, s, p, 1, 0, a.conv1d, ws, save_xd=True, offset_dev=rng_dev)
        bind.linear_bwd(dy, xd if xd is not None else x, W, A, B, H, keep, s, p, a.conv1d, ws, need_dx=True, dA=dA0,
                        dB=dB0, x_is_dropped=xd is not None)
        if exch is not None:
            exch.allreduce_avg_(bkt0)
        rng_dev.add_(4)

    if not a.no_graph and (exch is not None or not dist_on):
        try:
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(3):
                    graph_step()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                graph_step()
        except Exception as e:  # noqa: BLE001
            print(f"[bench] CUDA-graph capture failed ({e!r}); running eagerly", file=sys.stderr)
            graph = None
            if dist_on:
                raise  # ranks must not diverge between graph and eager paths

    run_step = (lambda i: graph.replay()) if graph is not None else step
    for i in range(a.warmup):
        run_step(i)
    drain()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(torch.cuda.current_device(), enabled=rank == 0) as clk:
        barrier()
        e0.record()
        for i in range(a.steps):
            run_step(i)
        drain()  # every all-reduce is inside the timed region
        e1.record()
        barrier()
    ms = e0.elapsed_time(e1)
    if dist_on:
        t = torch.tensor([ms], device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_step = ms / a.steps
    value = world * M / (ms_step * 1e-3)

    # ---- dominant kernel alone (tcgen05 GEMM of the forward), live, CUDA events on this stream ----
    try:
        Y, H, keep = bind.linear_fwd(x, W, b, A, B, s, p, 1, 0, a.conv1d, ws)  # leaves valid packed operands
        bind.set_phase_mask(bind.PHASE_GEMM)
        Ybuf = torch.empty_like(Y)
        n_rep = max(20, a.steps)
        for _ in range(5):
            bind.lib().lora_linear_fwd(x.data_ptr(), W.data_ptr(), b.data_ptr(), A.data_ptr(), B.data_ptr(),
                                       Ybuf.data_ptr(), H.data_ptr(), keep.data_ptr() if keep is not None else None,
                                       None, M, K, N, r, s, p, 1, 0,
                                       bind.dtype_code(dtype), int(a.conv1d), ws.data_ptr(), ws.numel(),
                                       torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(n_rep):
            bind.lib().lora_linear_fwd(x.data_ptr(), W.data_ptr(), b.data_ptr(), A.data_ptr(), B.data_ptr(),
                                       Ybuf.data_ptr(), H.data_ptr(), keep.data_ptr() if keep is not None else None,
                                       None, M, K, N, r, s, p, 1, 0,
                                       bind.dtype_code(dtype), int(a.conv1d), ws.data_ptr(), ws.numel(),
                                       torch.cuda.current_stream().cuda_stream)
        g1.record()
        torch.cuda.synchronize()
        gemm_ms = g0.elapsed_time(g1) / n_rep
    finally:
        bind.set_phase_mask(bind.PHASE_ALL)
    burst, sustained, src = measured_peaks()
    achieved = flops_gemm_launch(M, K, N, r) / (gemm_ms * 1e-3) / 1e12
    traffic = None
    tpath = os.path.join(REPO_ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(tpath):
        with open(tpath) as f:
            traffic = json.load(f).get(f"{M}x{K}x{N}")
    roofline = {"kernel": "lb::gemm::gemm_lora_kernel (forward launch)", "bound": "tensor", "achieved": achieved,
                "peak": burst, "unit": "TFLOP/s", "frac": achieved / burst, "traffic": traffic,
                "peak_source": f"{src} burst bf16 (MEASURED_PEAKS.json)", "kernel_ms": gemm_ms,
                "step_tflops": flops_fwd_bwd(M, K, N, r) / (ms_step * 1e-3) / 1e12,
                "step_frac_of_sustained": flops_fwd_bwd(M, K, N, r) / (ms_step * 1e-3) / 1e12 / sustained}

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
        esz = 2
        e2e = {"value": world * M / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": (M * K + M * N) * esz,
               "d2h_bytes_per_step": (M * N + M * K) * esz + (r * K + N * r) * 4, "ms_per_step": e2e_ms}

    # launches per step: fwd pack, rowproj, gemm; bwd pack, rowproj, gemm, colproj (dB+dA, reduction folded in)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": a.dtype, "data": "synthetic",
        "config": {"workload": workload_name(a), "M": M, "K": K, "N": N, "r": r, "lora_alpha": 16, "dropout": p,
                   "w_layout": "KN" if a.conv1d else "NK", "parallelism": f"dp{world}",
                   "l2_policy": "inputs larger than L2 (X, dY, Y, dX, W = %.0f MB per step vs 126 MB L2)" %
                                ((3 * M * K + 2 * M * N + K * N) * 2 / 1e6 if K == N else
                                 (2 * M * K + 2 * M * N + K * N) * 2 / 1e6),
                   "launch": "one CUDA-graph replay per step" if graph is not None else "eager",
                   "exchange": ("none" if not dist_on else
                                "in-stream peer-memory all-reduce(avg) of the flat adapter-grad bucket every step (csrc/p2p.cuh)"
                                if exch is not None else
                                "NCCL all-reduce(avg) of the flat adapter-grad bucket every step, overlapped with the next step's compute (two buckets)")},
        "clocks": clk.summary(), "e2e": e2e, "gpu_launches": (7 + (1 if exch is not None else 0)) * a.steps, "roofline": roofline,
    }
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
            f"{'stock torch lm_head + fp32 CE' if a.stock_head else 'lm_head on the tcgen05 GEMM + fused CE'}")


def run_model(a, pkg, rank, world, device):
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
        pkg.frozen.accelerate_lm_head(pm)
        pkg.frozen.accelerate_frozen_neighbours(pm)
        loss_fn = pkg.frozen.fused_causal_lm_loss
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
    e2e_ms, _ = timed(step_e2e, max(3, min(a.steps, 10)), 2)
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
        "gpu_launches": a.steps * (fwd_calls * 3 + n_lora * 4),
    }
    return line


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
        dist.init_process_group("nccl", device_id=device)
    try:
        if a.workload != "layer":
            line = run_model(a, pkg, rank, world, device)
            if rank == 0:
                print(json.dumps(line), flush=True)
            return
        if a.sweep:
            shapes = [(768, 2304, True), (768, 768, True), (3072, 768, True), (2560, 2560, False),
                      (2560, 10240, False), (10240, 2560, False)]
            for (K, N, conv) in shapes:
                for r in (8, 16, 64):
                    for M in (4096, 16384, 65536):
                        a.k, a.n, a.conv1d, a.r, a.tokens = K, N, conv, r, M
                        line = run_point(a, pkg, rank, world, device, with_e2e=False)
                        if rank == 0:
                            print(json.dumps({k: line[k] for k in ('value', 'ms_per_step', 'config', 'roofline')}), flush=True)
            return
        line = run_point(a, pkg, rank, world, device)
        if rank == 0:
            if world == 1 and not a.no_cpu_baseline:
                cores = use_all_host_threads()
                t = cpu_step_time(a, a.cpu_tokens, 5, 1)
                line["cpu_baseline"] = {"value": a.cpu_tokens / t, "unit": UNIT, "cores": cores, "kind": "port",
                                        "sample": f"{a.cpu_tokens} tokens/step x 5 steps, same layer shape, "
                                                  "oracle/lora_ref.py (torch fp32 CPU)"}
            print(json.dumps(line), flush=True)
    finally:
        if world > 1:
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
