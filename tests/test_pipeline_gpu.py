"""Multi-GPU: N NCCL pipeline stages reproduce the single-GPU tokens (needs >= 2 GPUs;
skipped on a 1-GPU box — the schedule itself is covered on CPU by test_pipeline_cpu.py)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("vp", [0, 1])
@pytest.mark.parametrize("world", [2, 4])
def test_nccl_pipeline_equals_single_engine(cuda, world, vp):
    """vp=1: lm_head split by vocabulary rows over the stages (VocabParallelRunner)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, have {torch.cuda.device_count()}")
    os.environ["DLI_PIPE_VP"] = str(vp)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29500 + world + 10 * vp), os.path.join(ROOT, "tests", "pipeline_check.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0 and "PIPELINE_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]


@pytest.mark.parametrize("world,vp,chunks", [(2, 0, 3), (2, 1, 4), (4, 1, 3)])
def test_nccl_pipeline_with_chunked_prefill(cuda, world, vp, chunks):
    """Prompts fed in chunks, the stages working on different chunks at once: tokens equal the
    single engine's with the same chunking."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, have {torch.cuda.device_count()}")
    env = dict(os.environ, DLI_PIPE_VP=str(vp), DLI_CHECK_CHUNKS=str(chunks))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29560 + world + 10 * vp), os.path.join(ROOT, "tests", "pipeline_check.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert p.returncode == 0 and "PIPELINE_OK" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]
