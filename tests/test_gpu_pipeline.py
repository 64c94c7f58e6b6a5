"""Pipeline-parallel LoRA step on the GPUs of one box (NCCL send/recv, this library's fused layers): what
tests/test_pipeline_cpu.py checks on gloo with the oracle's layer, here on the device path -- a pipeline is a
re-ordering, so loss, clip norm and every adapter gradient must equal those of the unsplit model run on one GPU
over the same micro-batches (fp16, BASELINE.json configs[4]; LoRA dropout off: split and unsplit models draw from
different generators).  Runs 2, 4 and 8 stages as far as the box has GPUs; skipped with fewer than two.
Reference intent: finetune_lora_opt_pp.py:156 (split), :121-153 (stage wrappers), :213-224 (train_batch)."""
import datetime
import os
import socket
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TARGETS = ["q_proj", "k_proj", "v_proj", "out_proj", "fc1", "fc2"]  # finetune_lora_opt_pp.py:109
N_LAYERS, N_MICRO = 8, 6


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _model(pkg, dev):
    from transformers import OPTConfig, OPTForCausalLM

    torch.manual_seed(0)
    m = OPTForCausalLM(OPTConfig(hidden_size=256, ffn_dim=1024, num_hidden_layers=N_LAYERS, num_attention_heads=4,
                                 word_embed_proj_dim=256, vocab_size=512, max_position_embeddings=128, dropout=0.0,
                                 attention_dropout=0.0, do_layer_norm_before=True, activation_function="relu"))
    m.config.use_cache = False
    pm = pkg.get_peft_model(m.to(torch.float16), pkg.LoraConfig(r=8, lora_alpha=16, target_modules=TARGETS, bias="none",
                                                               task_type=pkg.TaskType.CAUSAL_LM))
    g = torch.Generator().manual_seed(1)
    with torch.no_grad():
        for n, p in pm.named_parameters():
            if "lora_B" in n:
                p.copy_(torch.randn(p.shape, generator=g) * 0.05)
    return pm.to(dev)


def _worker(rank, world, port, q, overlap):
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
        batches = pkg.train.synthetic_batches(N_MICRO, batch=2, seq=64, vocab=512, pad_id=1, seed=77, device=dev)
        pm = _model(pkg, dev)
        stage = pkg.pipeline.OPTStage(pm.base_model.model, rank, world)
        eng = pkg.pipeline.PipelineEngine(stage, lr=1e-3, max_grad_norm=1.0, overlap=overlap)
        eng.zero_grad()
        loss = eng.train_batch(batches)
        torch.cuda.synchronize()
        counts = pkg.pipeline.split_layers(N_LAYERS, world)
        lo = sum(counts[:rank])
        mine = {}
        for n, p in stage.named_parameters():
            if p.requires_grad:
                local, rest = n.split(".", 2)[1:]
                mine[f"base_model.model.model.decoder.layers.{lo + int(local)}.{rest}"] = p.grad.float().cpu()
        shards = [None] * world
        dist.gather_object(mine, shards if rank == 0 else None, dst=0)
        info = [float(loss) if loss is not None else None, float(eng.last_grad_norm)]
        infos = [None] * world
        dist.gather_object(info, infos if rank == 0 else None, dst=0)
        if rank == 0:
            # the unsplit model on this GPU, same micro-batches, accumulate, clip
            ref = _model(pkg, dev)
            ref.train()
            tot = 0.0
            for b in batches:
                out = ref(input_ids=b["input_ids"], attention_mask=b["attention_mask"]).logits
                l = pkg.train.causal_lm_loss(out, b["labels"]) / N_MICRO
                l.backward()
                tot += float(l)
            params = [p for p in ref.parameters() if p.requires_grad]
            norm = float(torch.nn.utils.clip_grad_norm_(params, 1.0))
            want = {n: p.grad.float().cpu() for n, p in ref.named_parameters() if p.requires_grad}
            got = {}
            for s_ in shards:
                assert not (set(got) & set(s_)), "an adapter owned by two stages"
                got.update(s_)
            assert set(got) == set(want), (len(got), len(want))
            pipe_loss = infos[-1][0]
            assert abs(pipe_loss - tot) < 2e-3 * abs(tot), (pipe_loss, tot)
            assert all(abs(i[1] - norm) < 2e-2 * norm for i in infos), (infos, norm)
            worst = max(float((got[n] - want[n]).abs().max() / want[n].abs().max().clamp_min(1e-20)) for n in want)
            assert worst < 2e-2, worst
        dist.barrier()
        dist.destroy_process_group()
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        import traceback

        q.put((rank, "FAIL: " + repr(e) + "\n" + traceback.format_exc()))


@pytest.mark.timeout(900)
@pytest.mark.parametrize("world,overlap", [(2, True), (2, False), (4, True), (8, True)])
def test_pipeline_on_gpus_equals_unsplit_model(world, overlap):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs on one box")
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q, overlap)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in procs]
    for p in procs:
        p.join(60)
    assert all(r[1] == "ok" for r in res), res
