"""bench.py's contract on a machine without a GPU: the reference arm prints ONE JSON line with the driver's keys
(also under a torchrun-style environment, where only rank 0 may work), and the product arm refuses to run."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = {"impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
        "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"}


def _run(args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True,
                          cwd=ROOT, env=e, timeout=600)


def test_reference_arm_line():
    res = _run(["--impl", "reference", "--steps", "1", "--warmup", "1", "--cpu-tokens", "256", "--k", "256", "--n", "256"])
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [l for l in res.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert KEYS <= set(d) and d["impl"] == "reference" and d["value"] > 0 and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["metric"] == "lora_linear_fwd_bwd_tokens_per_sec" and d["unit"] == "tokens/s"


def test_reference_arm_other_ranks_do_no_work():
    res = _run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1", "--cpu-tokens", "256"],
               env={"RANK": "1", "LOCAL_RANK": "1", "WORLD_SIZE": "2", "OMP_NUM_THREADS": "1"})
    assert res.returncode == 0 and not [l for l in res.stdout.splitlines() if l.startswith("{")]


@pytest.mark.skipif(__import__("torch").cuda.is_available(), reason="checks the no-GPU refusal")
def test_product_arm_refuses_without_gpu():
    res = _run(["--steps", "1", "--warmup", "3"])
    assert res.returncode != 0 and "no CPU fallback" in (res.stderr + res.stdout)
