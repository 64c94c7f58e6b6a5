"""Multi-GPU diagnostic (torchrun): where does the layer step's time go when N processes run at once?
Per variant: device ms/step (max over ranks) and host enqueue us/step (max over ranks)."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import __graft_entry__ as g

pkg = g.load_package()
rank, world, local = pkg.train.init_distributed()
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
bind = pkg.binding
M, K, N, r, p, s = 16384, 2560, 2560, 8, 0.05, 2.0
gen = torch.Generator().manual_seed(rank)
x = torch.randn(M, K, generator=gen).to(dev, torch.bfloat16)
W = (torch.randn(N, K, generator=gen) * 0.02).to(dev, torch.bfloat16)
b = torch.randn(N, generator=gen).to(dev, torch.bfloat16)
A = (torch.randn(r, K, generator=gen) * 0.02).to(dev)
B = (torch.randn(N, r, generator=gen) * 0.02).to(dev)
dy = torch.randn(M, N, generator=gen).to(dev, torch.bfloat16)
ws = torch.zeros(bind.workspace_bytes(M, K, N, r), dtype=torch.uint8, device=dev)
packed = torch.empty(bind.packed_bytes(K, N, r), dtype=torch.uint8, device=dev)
bkt = torch.empty(r * K + N * r, device=dev)
dA, dB = bkt[: r * K].view(r, K), bkt[r * K:].view(N, r)
exch = pkg.comm.PeerBucketExchange(bkt.numel()) if world > 1 else None
rng = torch.zeros(1, dtype=torch.int64, device=dev)


def step(use_exch):
    Y, H, keep, _ = bind.linear_fwd(x, W, b, A, B, s, p, 1, 0, False, packed, offset_dev=rng)
    bind.linear_bwd(dy, x, W, packed, H, keep, r, s, p, False, ws, need_dx=True, dA=dA, dB=dB)
    if use_exch and exch is not None:
        exch.allreduce_avg_(bkt)
    rng.add_(4)


def measure(name, fn, steps):
    for _ in range(5):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    host_us = (time.perf_counter() - t0) / steps * 1e6
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps, host_us], device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(json.dumps({"variant": name, "world": world, "steps": steps, "ms_per_step": round(float(t[0]), 4),
                          "host_enqueue_us_per_step_max": round(float(t[1]), 1)}), flush=True)


def capture(use_exch):
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            step(use_exch)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        step(use_exch)
    return gr


for steps in (10, 100):
    measure("eager, no exchange", lambda: step(False), steps)
    measure("eager, p2p exchange", lambda: step(True), steps)
g0 = capture(False)
g1 = capture(True)
for steps in (10, 100):
    measure("graph, no exchange", g0.replay, steps)
    measure("graph, p2p exchange", g1.replay, steps)
if exch is not None:
    measure("exchange kernel alone (164 KB)", lambda: exch.allreduce_avg_(bkt), 100)
    big = torch.randn(1 << 20, device=dev)
    ex2 = pkg.comm.PeerBucketExchange(big.numel())
    measure("exchange alone, 4 MB one-shot", lambda: ex2.allreduce_avg_(big, 1), 50)
    measure("exchange alone, 4 MB two-shot", lambda: ex2.allreduce_avg_(big, 2), 50)
    measure("NCCL all_reduce alone, 164 KB", lambda: dist.all_reduce(bkt, op=dist.ReduceOp.AVG), 100)
    measure("NCCL all_reduce alone, 4 MB", lambda: dist.all_reduce(big, op=dist.ReduceOp.AVG), 50)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
