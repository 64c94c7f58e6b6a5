"""The in-stream peer-memory exchange of the adapter-gradient bucket (csrc/p2p.cuh) on >= 2 GPUs of one
box: against the fp64 mean of the per-rank inputs (which every rank can regenerate), bit-identical across
ranks, many calls in a row (parity reuse), one-shot / two-shot / ragged sizes, and as a DDP comm hook
against torch's own NCCL all-reduce.  Skipped where only one GPU is visible."""
import datetime
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    try:
        sys.path.insert(0, ROOT)
        import torch.distributed as dist
        import __graft_entry__ as g

        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                          LOCAL_RANK=str(rank))
        pkg = g.load_package()
        torch.cuda.set_device(rank)
        dev = torch.device("cuda", rank)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev, timeout=datetime.timedelta(seconds=120))
        ex = pkg.comm.PeerBucketExchange(1 << 20)

        def data(r, step, n):
            return torch.randn(n, generator=torch.Generator().manual_seed(1000 * step + r), dtype=torch.float32)

        step = 0
        for n, mode in [(40960, 0), (1003, 0), (4, 1), (1 << 20, 2), (1 << 20, 1), (4096, 2), (262144 + 8, 0)]:
            for rep in range(5):  # consecutive calls reuse both parities of the exchange buffers
                step += 1
                t = data(rank, step, n).to(dev)
                ex.allreduce_avg_(t, mode)
                want = sum(data(r, step, n).double() for r in range(world)) / world
                err = float((t.cpu().double() - want).abs().max())
                assert err < 1e-6, (n, mode, rep, err)
                gathered = [torch.empty_like(t) for _ in range(world)]
                dist.all_gather(gathered, t)
                assert all(torch.equal(gathered[0], x) for x in gathered), "ranks disagree bitwise"

        # as the DDP communication hook
        torch.manual_seed(0)
        net = torch.nn.Sequential(torch.nn.Linear(64, 64), torch.nn.ReLU(), torch.nn.Linear(64, 8)).to(dev)
        x = torch.randn(16, 64, generator=torch.Generator().manual_seed(50 + rank)).to(dev)
        grads = {}
        for exch in ("nccl", "p2p"):
            net.zero_grad(set_to_none=True)
            ddp = pkg.train.wrap_ddp(net, dev, exchange=exch)
            for _ in range(3):
                net.zero_grad(set_to_none=True)
                ddp(x).square().mean().backward()
            grads[exch] = [p.grad.clone() for p in net.parameters()]
            pkg.train.close_exchange(ddp)  # collective: own buffer and the peers' IPC mappings
            del ddp
        for a, b in zip(grads["nccl"], grads["p2p"]):
            assert torch.allclose(a, b, rtol=1e-5, atol=1e-7)
        torch.cuda.synchronize()
        ex.close()
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback

        q.put((rank, "FAIL: " + repr(e) + "\n" + traceback.format_exc()))


@pytest.mark.timeout(600)
def test_peer_bucket_exchange_two_or_more_gpus():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs on one box")
    import torch.multiprocessing as mp

    world = min(n, 8)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, 29611, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(60)
    assert all(r[1] == "ok" for r in res), res
