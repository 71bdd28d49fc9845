"""bench.py contract checks that need no GPU: the reference arm prints one JSON line with the
agreed keys, non-zero ranks of a torchrun launch stay silent, and the product arm refuses to
run without a CUDA device (there is no CPU fallback to time by accident)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, env_extra=None, timeout=600):
    env = dict(os.environ)
    env.pop("RANK", None), env.pop("WORLD_SIZE", None), env.pop("LOCAL_RANK", None)
    env.update(env_extra or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, env=env,
                          capture_output=True, text=True, timeout=timeout)


def test_reference_arm_prints_one_contract_line():
    p = _run(["--impl", "reference", "--model", "gpt2", "--steps", "2", "--warmup", "1"])
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "decode_tokens_per_s" and d["unit"] == "tokens/s"
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["n_gpus"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0
    assert "workload" in d["config"] and "model" not in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_uses_all_cores_under_torchrun_and_other_ranks_are_silent():
    # torchrun exports OMP_NUM_THREADS=1; rank 0 must still use every core it may run on
    p0 = _run(["--impl", "reference", "--model", "gpt2", "--gpus", "2", "--steps", "1", "--warmup", "0"],
              {"RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0", "OMP_NUM_THREADS": "1"})
    assert p0.returncode == 0, p0.stderr[-2000:]
    d = json.loads([ln for ln in p0.stdout.splitlines() if ln.startswith("{")][0])
    ncpu = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    assert d["cpu_baseline"]["cores"] == ncpu and d["n_gpus"] == 2
    p1 = _run(["--impl", "reference", "--model", "gpt2", "--gpus", "2", "--steps", "1", "--warmup", "0"],
              {"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert p1.returncode == 0 and not [ln for ln in p1.stdout.splitlines() if ln.startswith("{")]


def test_product_arm_needs_a_gpu():
    import torch

    if torch.cuda.is_available():
        import pytest

        pytest.skip("a GPU is present: the product arm would run")
    p = _run(["--model", "gpt2", "--steps", "1", "--warmup", "1", "--no-cpu-baseline", "--no-extra"])
    assert p.returncode != 0
    assert not [ln for ln in p.stdout.splitlines() if ln.startswith('{"metric"')]
    assert "no CPU path" in p.stderr or "CUDA" in p.stderr
