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


@pytest.mark.parametrize("world,vp,hop", [(2, 1, 2), (2, 1, 1), (2, 0, 2), (4, 1, 2), (4, 0, 1)])
def test_nccl_mailbox_pipeline_equals_single_engine(cuda, world, vp, hop):
    """Decode over per-tick mailboxes (one NCCL all-gather per tick; hop 2 = on its own stream under the next tick's
    kernels): tokens equal the single engine's."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, have {torch.cuda.device_count()}")
    env = dict(os.environ, DLI_PIPE_VP=str(vp), DLI_CHECK_MAIL=str(hop))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29620 + world + 10 * vp + hop), os.path.join(ROOT, "tests", "pipeline_check.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
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


def test_worker_group_serves_a_sharded_model_over_http(cuda, tmp_path):
    """The deployment of the north star end to end: two reference-style workers (HTTP, one per GPU) in one
    process group; /load_shard registers shard i on worker i, /inference on worker 0 runs the NCCL pipeline.
    Same text as one worker holding both shards."""
    import json
    import signal
    import socket
    import time

    import requests

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from golden_util import load_case, spec_kwargs
    from werkzeug.test import Client

    from dli_b200.engine import ModelSpec
    from dli_b200.sharding import write_shards
    from dli_b200.worker import Worker
    from oracle import decoder_ref as R

    c = load_case("llama_tiny")
    cfg = spec_kwargs(c["config"])
    spec = ModelSpec.from_hf_config(cfg)
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"])
    paths = write_shards(sd, cfg, "org/tiny", 2, str(tmp_path))
    full = write_shards(sd, cfg, "org/tiny-full", 1, str(tmp_path))[0]
    prompt = " ".join(str(int(i)) for i in c["prompts"][0])
    T = c["prompts"].shape[1]
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    env = dict(os.environ, PYTHONPATH=ROOT, USE_GPU="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29611", "-m", "dli_b200.pipeline_worker", "--host", "127.0.0.1", "--port", str(port)]
    p = subprocess.Popen(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                         start_new_session=True)
    try:
        for r in range(2):  # both workers answer /health
            for _ in range(240):
                try:
                    if requests.get(f"http://127.0.0.1:{port + r}/health", timeout=2).status_code == 200:
                        break
                except requests.RequestException:
                    time.sleep(0.5)
            else:
                raise AssertionError("worker group did not come up")
        for r in range(2):
            resp = requests.post(f"http://127.0.0.1:{port + r}/load_shard",
                                 json={"model_name": "org/tiny", "shard_id": r, "shard_path": paths[r]}, timeout=60)
            assert resp.status_code == 200, resp.text
        req = {"model_name": "org/tiny", "prompt": prompt, "shard_ids": [0, 1], "max_length": T + 20, "do_sample": False}
        resp = requests.post(f"http://127.0.0.1:{port}/inference", json=req, timeout=300)
        assert resp.status_code == 200, resp.text
        got = resp.json()["result"]
        again = requests.post(f"http://127.0.0.1:{port}/inference", json=req, timeout=300).json()["result"]
        sampled = [requests.post(f"http://127.0.0.1:{port}/inference", json=dict(req, do_sample=True, seed=4),
                                 timeout=300).json()["result"] for _ in range(2)]
        # a shard that is not there fails the request for everybody, and the group keeps serving
        bad = requests.post(f"http://127.0.0.1:{port}/inference", json=dict(req, model_name="org/none"), timeout=60)
        assert bad.status_code == 500
        too_short = requests.post(f"http://127.0.0.1:{port}/inference", json=dict(req, max_length=T), timeout=60)
        assert too_short.status_code == 500 and "Input length of input_ids" in too_short.json()["message"]
        assert requests.post(f"http://127.0.0.1:{port}/inference", json=req, timeout=300).json()["result"] == got
        # what the unchanged reference master sends: only the FIRST node's shard ids (views.py:337-340) — the group
        # still runs the whole pipeline; and a request that lands on the other node is handed to the pipeline head
        only_first = requests.post(f"http://127.0.0.1:{port}/inference", json=dict(req, shard_ids=[0]), timeout=300)
        assert only_first.status_code == 200 and only_first.json()["result"] == got, only_first.text
        on_other = requests.post(f"http://127.0.0.1:{port + 1}/inference", json=dict(req, shard_ids=[1]), timeout=300)
        assert on_other.status_code == 200 and on_other.json()["result"] == got, on_other.text
        # concurrent requests share the pipeline's slots (the reference serialises them, worker/Dockerfile:45):
        # every one gets exactly the text it gets alone
        from concurrent.futures import ThreadPoolExecutor

        others = [dict(req, prompt=" ".join(str(int(i)) for i in c["prompts"][k % c["prompts"].shape[0]].roll(k)),
                       max_length=T + 12 + 3 * k, do_sample=bool(k % 2), seed=k) for k in range(5)]
        alone = [requests.post(f"http://127.0.0.1:{port}/inference", json=o, timeout=300).json()["result"] for o in others]
        with ThreadPoolExecutor(max_workers=5) as ex:
            together = list(ex.map(lambda o: requests.post(f"http://127.0.0.1:{port}/inference", json=o, timeout=300).json(), others))
        assert [t.get("result") for t in together] == alone, together
        assert len(alone[0].split()) == T + 12 and len(alone[4].split()) == T + 24
        # worker 1 is a complete worker too: an unsharded model on ITS GPU gives the same text
        solo = requests.post(f"http://127.0.0.1:{port + 1}/inference",
                             json={"model_name": full, "prompt": prompt, "max_length": T + 20, "do_sample": False},
                             timeout=300)
        assert solo.status_code == 200 and solo.json()["result"] == got, solo.text
        # "stream": true — text so far after every block of rounds, then the plain answer
        with requests.post(f"http://127.0.0.1:{port}/inference", json=dict(req, max_length=T + 40, stream=True),
                           timeout=300, stream=True) as sr:
            lines = [json.loads(ln) for ln in sr.iter_lines() if ln]
        long_plain = requests.post(f"http://127.0.0.1:{port}/inference", json=dict(req, max_length=T + 40), timeout=300).json()
        assert lines[-1]["http_status"] == 200 and lines[-1]["result"] == long_plain["result"], lines[-1]
        parts = [ln["result"] for ln in lines[:-1]]
        assert len(parts) >= 2 and all(b.startswith(a) for a, b in zip(parts, parts[1:] + [lines[-1]["result"]]))
        # /unload_model on ONE node while the group has the model resident: the next request fails with the reference's
        # message for everybody, nothing crashes, and after the shard is registered again the group rebuilds and serves
        assert requests.post(f"http://127.0.0.1:{port + 1}/unload_model", json={"model_name": "org/tiny"}, timeout=60).status_code == 200
        gone = requests.post(f"http://127.0.0.1:{port}/inference", json=req, timeout=120)
        assert gone.status_code == 500 and "Shard 1 of model org/tiny is not loaded" in gone.json()["message"], gone.text
        assert requests.post(f"http://127.0.0.1:{port + 1}/load_shard",
                             json={"model_name": "org/tiny", "shard_id": 1, "shard_path": paths[1]}, timeout=60).status_code == 200
        back = requests.post(f"http://127.0.0.1:{port}/inference", json=req, timeout=300)
        assert back.status_code == 200 and back.json()["result"] == got, back.text
    finally:
        os.killpg(p.pid, signal.SIGTERM)
        try:
            out = p.communicate(timeout=30)[0]
        except subprocess.TimeoutExpired:
            os.killpg(p.pid, signal.SIGKILL)
            out = p.communicate()[0]
    # one worker holding both shards (LocalChain on one GPU): same kernels, same text
    w = Worker({"MODEL_CACHE_DIR": str(tmp_path)})
    cl = Client(w)
    for r in range(2):
        assert cl.post("/load_shard", json={"model_name": "org/tiny", "shard_id": r, "shard_path": paths[r]}).status_code == 200
    want = json.loads(cl.post("/inference", json=req).get_data(as_text=True))["result"]
    assert got == want and again == got, (got, want, out[-2000:])
    assert sampled[0] == sampled[1] and len(sampled[0].split()) == T + 20
    # The group splits the head over the stages and draws on stage 0 from the logits row gathered in the mails, keyed
    # by (seed, token position) like the continuous batcher: same draw, same text as a batching worker that holds the
    # whole model and the whole head.
    wb = Worker({"MODEL_CACHE_DIR": str(tmp_path), "DLI_BATCHING": "1", "DLI_MAX_BATCH": "2"})
    cb = Client(wb)
    try:
        want_sampled = json.loads(cb.post("/inference", json={"model_name": full, "prompt": prompt, "max_length": T + 20,
                                                              "do_sample": True, "seed": 4}).get_data(as_text=True))
        assert want_sampled.get("result") == sampled[0], (sampled[0], want_sampled)
    finally:
        cb.post("/unload_model", json={"model_name": full})

