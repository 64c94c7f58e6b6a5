"""Data-parallel harness on CPU: world_size-2 gloo run of the Trainer + DDP wrapper with the
oracle's LoRA layer (BASELINE.json configs[0]: distilgpt2-style LoRA r=8 on c_attn/c_proj, DDP on
CPU/gloo).  Checks what the reference relies on (finetune_lora_distilgpt2.py:139-144): only the
adapter parameters enter the reducer, ranks end with identical adapters, and the 2-rank run equals
a single-process run over the concatenated batch (DDP averages gradients)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REPO_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _model():
    from transformers import GPT2Config, GPT2LMHeadModel

    from oracle import lora_ref

    torch.manual_seed(0)
    m = GPT2LMHeadModel(GPT2Config(n_layer=2, n_embd=64, n_head=2, vocab_size=128, n_positions=32, resid_pdrop=0.0,
                                   embd_pdrop=0.0, attn_pdrop=0.0))
    pm = lora_ref.ref_get_peft_model(m, lora_ref.RefLoraConfig(r=8, lora_alpha=16, target_modules=["c_attn", "c_proj"]))
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for n, p in pm.named_parameters():
            if "lora_B" in n:
                p.copy_(torch.randn(p.shape, generator=g) * 0.02)
    return pm


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, REPO_ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.set_num_threads(1)
    import __graft_entry__ as g

    pkg = g.load_package()
    r, w, _ = pkg.train.init_distributed("gloo")
    assert (r, w) == (rank, world)
    pm = _model()
    ddp = pkg.train.wrap_ddp(pm)
    n_reduced = sum(p.numel() for p in ddp.parameters() if p.requires_grad)
    batches = pkg.train.synthetic_batches(8, batch=2, seq=16, vocab=128, pad_id=127, seed=1234 + rank)
    tr = pkg.train.Trainer(ddp, pkg.train.TrainConfig(lr=1e-2, accum=2))
    losses = [float(l) for l in tr.train(batches)]
    sd = {n: p.detach().clone() for n, p in pm.named_parameters() if p.requires_grad}
    torch.save({"losses": losses, "sd": sd, "n_reduced": n_reduced}, os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_ddp_gloo_world2_matches_single_process(tmp_path, pkg):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = torch.load(tmp_path / "rank0.pt")
    r1 = torch.load(tmp_path / "rank1.pt")
    # only adapters are trainable / reduced: 2 layers x (c_attn 64->192, attn.c_proj 64->64, mlp.c_proj 256->64)
    expect = 2 * 8 * ((64 + 192) + (64 + 64) + (256 + 64))
    assert r0["n_reduced"] == expect
    for n in r0["sd"]:
        assert torch.equal(r0["sd"][n], r1["sd"][n]), n  # ranks stay in lock-step
    # single process over the concatenated per-step batch == the 2-rank DDP run (mean of equal-size means)
    sys.path.insert(0, REPO_ROOT)
    pm = _model()
    b0 = pkg.train.synthetic_batches(8, batch=2, seq=16, vocab=128, pad_id=127, seed=1234)
    b1 = pkg.train.synthetic_batches(8, batch=2, seq=16, vocab=128, pad_id=127, seed=1235)
    cat = [{k: torch.cat([x[k], y[k]]) for k in x} for x, y in zip(b0, b1)]
    tr = pkg.train.Trainer(pm, pkg.train.TrainConfig(lr=1e-2, accum=2))
    tr.train(cat)
    for n, p in pm.named_parameters():
        if p.requires_grad:
            assert torch.allclose(p, r0["sd"][n], rtol=1e-4, atol=1e-6), n
