"""Pipeline-parallel harness on CPU (gloo, 2 and 3 stages) with the oracle's LoRA layer on a tiny
OPT: a pipeline is a re-ordering, so the loss and every adapter gradient/update must equal those
of the unsplit model accumulated over the same micro-batches ("P2 has no
executable semantics -- the oracle is the unsplit model")."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

REPO_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TARGETS = ["q_proj", "k_proj", "v_proj", "out_proj", "fc1", "fc2"]  # finetune_lora_opt_pp.py:109


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _model(n_layers):
    from transformers import OPTConfig, OPTForCausalLM

    from oracle import lora_ref

    torch.manual_seed(0)
    m = OPTForCausalLM(OPTConfig(hidden_size=32, ffn_dim=64, num_hidden_layers=n_layers, num_attention_heads=2,
                                 word_embed_proj_dim=32, vocab_size=64, max_position_embeddings=32, dropout=0.0,
                                 attention_dropout=0.0, do_layer_norm_before=True))
    pm = lora_ref.ref_get_peft_model(m, lora_ref.RefLoraConfig(r=4, lora_alpha=8, target_modules=TARGETS))
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for n, p in pm.named_parameters():
            if "lora_B" in n:
                p.copy_(torch.randn(p.shape, generator=g) * 0.05)
    return pm


def _batches(pkg, n):
    return pkg.train.synthetic_batches(n, batch=2, seq=12, vocab=64, pad_id=1, seed=77)


def _worker(rank, world, port, out_dir, n_layers, n_micro, overlap):
    sys.path.insert(0, REPO_ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    torch.set_num_threads(1)
    import __graft_entry__ as g

    pkg = g.load_package()
    pkg.train.init_distributed("gloo")
    pm = _model(n_layers)
    stage = pkg.pipeline.OPTStage(pm.base_model.model, rank, world)  # finetune_lora_opt_pp.py:116
    eng = pkg.pipeline.PipelineEngine(stage, lr=1e-2, max_grad_norm=1.0, overlap=overlap)
    losses = []
    for step in range(2):
        eng.zero_grad()
        loss = eng.train_batch(_batches(pkg, n_micro))
        losses.append(None if loss is None else float(loss))
    sd = {n: p.detach().clone() for n, p in pm.named_parameters() if p.requires_grad}
    owned = [n for n, p in stage.named_parameters() if p.requires_grad]
    torch.save({"losses": losses, "sd": sd, "owned": owned, "norm": float(eng.last_grad_norm)},
               os.path.join(out_dir, f"rank{rank}.pt"))
    cfg = pkg.LoraConfig(r=4, lora_alpha=8, target_modules=TARGETS, task_type=pkg.TaskType.CAUSAL_LM)
    pkg.pipeline.save_adapters(stage, os.path.join(out_dir, "adapters"), peft_config=cfg)  # finetune_lora_opt_pp.py:226-243
    dist.barrier()
    dist.destroy_process_group()


# overlap=True: sends and receives on per-direction communicators, nothing grouped; False: grouped blocking exchanges
@pytest.mark.parametrize("world,n_layers,n_micro,overlap", [(2, 4, 4, True), (3, 5, 2, False)])
def test_pipeline_equals_unsplit_model(tmp_path, pkg, world, n_layers, n_micro, overlap):
    assert pkg.pipeline.split_layers(32, 8) == [4] * 8 and pkg.pipeline.split_layers(5, 3) == [2, 2, 1]
    port = _free_port()
    mp.spawn(_worker, args=(world, port, str(tmp_path), n_layers, n_micro, overlap), nprocs=world, join=True)
    ranks = [torch.load(tmp_path / f"rank{r}.pt") for r in range(world)]
    # unsplit reference: same micro-batches, accumulate, clip, AdamW -- twice
    from oracle import lora_ref

    pm = _model(n_layers)
    params = [p for p in pm.parameters() if p.requires_grad]
    opt = torch.optim.AdamW(params, lr=1e-2, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0)
    ref_losses = []
    pm.train()
    for step in range(2):
        opt.zero_grad(set_to_none=True)
        tot = 0.0
        for b in _batches(pkg, n_micro):
            loss = lora_ref.causal_lm_loss(pm(input_ids=b["input_ids"], attention_mask=b["attention_mask"]).logits,
                                           b["labels"]) / n_micro
            loss.backward()
            tot += float(loss)
        norm = torch.nn.utils.clip_grad_norm_(params, 1.0)
        opt.step()
        ref_losses.append(tot)
    last = ranks[-1]
    assert all(abs(a - b) < 1e-5 for a, b in zip(last["losses"], ref_losses)), (last["losses"], ref_losses)
    assert abs(last["norm"] - float(norm)) < 1e-4 * max(1.0, float(norm))
    ref_sd = dict(pm.named_parameters())
    seen = set()
    # map stage-owned names (layers.<local idx>...) back to global names through the split
    counts = pkg.pipeline.split_layers(n_layers, world)
    for rank, r in enumerate(ranks):
        lo = sum(counts[:rank])
        for o in r["owned"]:
            assert o.startswith("layers."), o
            local, rest = o.split(".", 2)[1:]
            gname = f"base_model.model.model.decoder.layers.{lo + int(local)}.{rest}"
            assert torch.allclose(r["sd"][gname], ref_sd[gname], rtol=1e-4, atol=1e-6), gname
            seen.add(gname)
    assert seen == {n for n, p in pm.named_parameters() if p.requires_grad}
    # rank 0 wrote ONE peft-format directory holding every stage's adapters
    import json

    from safetensors.torch import load_file

    saved = load_file(str(tmp_path / "adapters" / "adapter_model.safetensors"))
    want = {n.replace(".default.", "."): p for n, p in pm.named_parameters() if p.requires_grad}
    assert set(saved) == set(want)
    for k, v in saved.items():
        assert torch.allclose(v, want[k].detach(), rtol=1e-4, atol=1e-6), k
    meta = json.load(open(tmp_path / "adapters" / "meta.json"))
    assert meta == {"split": counts, "stages": world}
    assert json.load(open(tmp_path / "adapters" / "adapter_config.json"))["r"] == 4
