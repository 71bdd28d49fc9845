"""HTTP contract of the drop-in worker, checked against the exchanges recorded from the
reference's own worker/app.py (tests/golden/worker_contract.json, oracle/gen_golden.py).
CPU part: every exchange that needs no model arithmetic.  GPU part (marked): /inference
through the CUDA engine, result equal to the oracle's greedy continuation."""
import json
import os

import pytest
import torch
from werkzeug.test import Client

from golden_util import GOLD, load_case, spec_kwargs

GOLDEN = json.load(open(os.path.join(GOLD, "worker_contract.json")))


def _client(env=None):
    from dli_b200.worker import Worker

    w = Worker(env={"MODEL_CACHE_DIR": "/nonexistent", **(env or {})})
    return Client(w), w


def _post(c, path, body, headers=None):
    r = c.post(path, json=body, headers=headers or {})
    return r.status_code, r.get_json()


def _golden(route, note):
    hits = [g for g in GOLDEN if g["route"] == route and g["note"] == note]
    assert len(hits) == 1, (route, note)
    return hits[0]


def test_error_paths_match_reference_byte_for_byte(tmp_path):
    c, _ = _client()
    for route, note, body in [
        ("/load_model", "missing model_name", {}),
        ("/load_shard", "missing shard_path", {"model_name": "tiny", "shard_id": 1}),
        ("/inference", "missing prompt", {"model_name": "tiny"}),
        ("/unload_model", "missing model_name", {}),
        ("/unload_model", "unloading nothing succeeds", {"model_name": "never-loaded"}),
    ]:
        g = _golden(route, note)
        assert g["request"] == body
        assert _post(c, route, body) == (g["status"], g["response"]), (route, note)


def test_unload_model_keeps_other_models_with_a_common_prefix():
    """The reference deletes exactly loaded_models[model_name] (worker/app.py:223-227); the per-shard engines this
    worker creates lazily are "<model>_<shard id>" entries and go too (Q7) — but "gpt2_medium" is not a shard of
    "gpt2"."""
    c, w = _client()

    class Fake:
        unloaded = False

        def unload(self):
            self.unloaded = True

    full, shard, other = Fake(), Fake(), Fake()
    w.loaded_models.update({"gpt2": full, "gpt2_1": shard, "gpt2_medium": other})
    w.loaded_tokenizers.update({"gpt2": object(), "gpt2_medium": object()})
    w.loaded_shards["gpt2"] = {1: {"path": "/x", "metadata": {}}}
    assert _post(c, "/unload_model", {"model_name": "gpt2"})[0] == 200
    assert full.unloaded and shard.unloaded and not other.unloaded
    assert list(w.loaded_models) == ["gpt2_medium"] and list(w.loaded_tokenizers) == ["gpt2_medium"]
    assert "gpt2" not in w.loaded_shards


def test_shard_registration_health_and_unload_match_reference(tmp_path):
    c, w = _client()
    shard_dir = tmp_path / "tiny" / "shard_1"
    shard_dir.mkdir(parents=True)
    meta = _golden("/health", "one shard registered")["response"]["loaded_shards"][0]["metadata"]
    json.dump(meta, open(shard_dir / "metadata.json", "w"))
    body = {"model_name": "tiny", "shard_id": 1, "shard_path": str(shard_dir)}
    g = _golden("/load_shard", "")
    assert _post(c, "/load_shard", body) == (g["status"], g["response"])
    g = _golden("/load_shard", "idempotent")
    assert _post(c, "/load_shard", body) == (g["status"], g["response"])
    # /health: same keys, same shard entry (path differs by machine)
    g = _golden("/health", "one shard registered")["response"]
    r = c.get("/health")
    h = r.get_json()
    assert r.status_code == 200 and set(h) == set(g) and set(h["resources"]) == set(g["resources"])
    assert h["status"] == "healthy" and h["loaded_models"] == g["loaded_models"]
    assert h["loaded_tokenizers"] == g["loaded_tokenizers"]
    (mine,), (theirs,) = h["loaded_shards"], g["loaded_shards"]
    assert {k: mine[k] for k in ("model_name", "shard_id", "metadata")} == \
           {k: theirs[k] for k in ("model_name", "shard_id", "metadata")}
    # inference on a shard that is not registered: same 500 and message as the reference
    g = _golden("/inference", "shard not loaded")
    assert _post(c, "/inference", g["request"]) == (g["status"], g["response"])
    g = _golden("/unload_model", "")
    assert _post(c, "/unload_model", {"model_name": "tiny"}) == (g["status"], g["response"])
    assert c.get("/health").get_json()["loaded_shards"] == []
    # metadata.json absent -> the reference's synthesised default (worker/app.py:181-189)
    bare = tmp_path / "bare"
    bare.mkdir()
    assert _post(c, "/load_shard", {"model_name": "m2", "shard_id": "a", "shard_path": str(bare)})[0] == 200
    assert w.loaded_shards["m2"]["a"]["metadata"] == {"model_name": "m2", "shard_id": "a", "num_shards": 1,
                                                     "start_layer": 0, "end_layer": 0, "total_layers": 1}


def test_shard_id_zero_is_accepted_deliberate_deviation(tmp_path):
    """The reference answers 400 here (falsy 0 in all([...])); a first shard
    must be registrable, so this worker accepts it."""
    g = _golden("/load_shard", "shard_id 0 is falsy (Q1)")
    assert g["status"] == 400
    c, _ = _client()
    code, body = _post(c, "/load_shard", {"model_name": "tiny", "shard_id": 0, "shard_path": str(tmp_path)})
    assert code == 200 and body["message"] == "Shard 0 of model tiny loaded successfully"
    assert _post(c, "/load_shard", {"model_name": "tiny", "shard_path": str(tmp_path)})[0] == 400


def test_auth_matches_reference():
    c, _ = _client({"AUTH_ENABLED": "1", "AUTH_KEY": "k"})
    g = _golden("/health", "AUTH_ENABLED=1, no header")
    r = c.get("/health")
    assert (r.status_code, r.get_json()) == (g["status"], g["response"])
    r = c.get("/health", headers={"Authorization": "Bearer k"})
    assert r.status_code == 200 and r.get_json()["status"] == "healthy"
    assert c.post("/inference", json={"model_name": "a", "prompt": "1"}).status_code == 401


def test_unknown_route_and_no_cpu_path():
    c, _ = _client({"USE_GPU": "0"})
    assert c.get("/nope").status_code == 404
    assert c.get("/inference").status_code == 405
    code, body = _post(c, "/load_model", {"model_name": "org/none"})
    assert code == 500 and body["message"].startswith("Failed to load model: ")


# ------------------------------------------------------------------ GPU: /inference end to end
@pytest.fixture(scope="module")
def shard_dirs(tmp_path_factory):
    from dli_b200.sharding import write_shards
    from oracle import decoder_ref as R

    from dli_b200.engine import ModelSpec

    c = load_case("llama_tiny")
    cfg = spec_kwargs(c["config"])
    spec = ModelSpec.from_hf_config(cfg)
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"])
    out = tmp_path_factory.mktemp("shards")
    paths = write_shards(sd, cfg, "org/tiny", 2, str(out))
    full = write_shards(sd, cfg, "org/tiny-full", 1, str(out))[0]
    return c, spec, sd, paths, full


@pytest.mark.gpu
def test_inference_sharded_and_unsharded_equal_oracle(cuda, shard_dirs):
    from oracle import decoder_ref as R

    case, spec, sd, paths, full = shard_dirs
    c, w = _client({"MODEL_CACHE_DIR": os.path.dirname(full)})
    ids = case["prompts"][0]
    prompt = " ".join(str(int(i)) for i in ids)
    T, n_new = len(ids), 6
    want = R.greedy_generate(ids[None], sd, R.ModelSpec(**spec.__dict__), T + n_new)
    eng_want = None
    # sharded: both layer ranges resident on this worker, ids 0 and 1
    for i, p in enumerate(paths):
        assert _post(c, "/load_shard", {"model_name": "org/tiny", "shard_id": i, "shard_path": p})[0] == 200
    code, body = _post(c, "/inference", {"model_name": "org/tiny", "prompt": prompt, "shard_ids": [0, 1],
                                         "max_length": T + n_new, "do_sample": False})
    assert code == 200 and set(body) == {"status", "result", "execution_time"} and body["status"] == "success"
    got = [int(t) for t in body["result"].split()]
    assert got[:T] == ids.tolist() and len(got) == T + n_new
    # unsharded path through /load_model (lazy, like the reference) gives the same tokens
    code, body2 = _post(c, "/inference", {"model_name": full, "prompt": prompt, "max_length": T + n_new,
                                          "do_sample": False})
    assert code == 200 and body2["result"] == body["result"]
    # against the oracle: identical over the margin-safe prefix of this fixture sequence
    noise = 2.0 * case["hf_err"].max()
    n_safe = int((case["margin"][:, 0] > 2.0 * noise).int().cumprod(0).sum())
    assert got[T:T + min(n_safe, n_new)] == want[0, T:T + min(n_safe, n_new)].tolist()
    # default request = the reference's sampling call; a seed makes it reproducible
    req = {"model_name": full, "prompt": prompt, "max_length": T + n_new, "seed": 5}
    (c1, b1), (c2, b2) = _post(c, "/inference", req), _post(c, "/inference", req)
    assert c1 == c2 == 200 and b1["result"] == b2["result"] and len(b1["result"].split()) == T + n_new
    # a shard subset that does not cover all layers is refused loudly
    code, body3 = _post(c, "/inference", {"model_name": "org/tiny", "prompt": prompt, "shard_ids": [1]})
    assert code == 500 and "not all of" in body3["message"]
    # timeout -> 408 with the reference's message
    g = _golden("/inference", "timeout -> 408")
    code, body4 = _post(c, "/inference", {"model_name": full, "prompt": prompt, "max_length": T + 4, "timeout": -1})
    assert (code, body4) == (g["status"], g["response"])
    # a prompt that already fills max_length: transformers refuses, the reference answers 500 with its message
    for req in ({"model_name": full, "prompt": prompt, "max_length": T},
                {"model_name": "org/tiny", "prompt": prompt, "shard_ids": [0, 1], "max_length": T - 2}):
        code, body5 = _post(c, "/inference", req)
        assert code == 500 and body5["status"] == "error"
        assert body5["message"].startswith(f"Inference failed: Input length of input_ids is {T}, but `max_length` is set to")
    # unload frees the lazily created per-shard engines too (Q7)
    assert _post(c, "/unload_model", {"model_name": "org/tiny"})[0] == 200
    assert not [k for k in w.loaded_models if k.startswith("org/tiny_")]


def test_stop_tokens_come_from_generation_config_first(tmp_path):
    """transformers takes eos/pad from generation_config.json when the checkpoint has one, else from
    config.json; an int or a list are both legal."""
    import json

    from dli_b200.worker import _stop_tokens

    cfg = {"eos_token_id": 2, "pad_token_id": None}
    assert _stop_tokens(str(tmp_path), cfg) == ((2,), None)
    assert _stop_tokens(str(tmp_path), {}) == ((), None)
    (tmp_path / "generation_config.json").write_text(json.dumps({"eos_token_id": [128001, 128009], "pad_token_id": 0}))
    assert _stop_tokens(str(tmp_path), cfg) == ((128001, 128009), 0)


# ------------------------------------------------------------------ streaming (an extension: "stream": true)
def _lines(resp):
    return [json.loads(ln) for ln in resp.get_data(as_text=True).splitlines() if ln.strip()]


def test_streaming_request_errors_keep_the_reference_shapes():
    """Without "stream" nothing changes; with it, malformed / unauthorised requests still get the reference's plain JSON
    and status, and a failure inside the request arrives as the last line with its status under "http_status"."""
    c, _ = _client({"USE_GPU": "0"})
    r = c.post("/inference", json={"model_name": "m", "stream": True})
    assert r.status_code == 400 and r.get_json() == {"status": "error", "message": "Model name and prompt are required"}
    ca, _ = _client({"AUTH_ENABLED": "1", "AUTH_KEY": "k"})
    r = ca.post("/inference", json={"model_name": "m", "prompt": "1 2", "stream": True})
    assert r.status_code == 401 and r.get_json()["message"] == "Unauthorized access"
    r = c.post("/inference", json={"model_name": "org/none", "prompt": "1 2", "stream": True})
    assert r.status_code == 200 and r.mimetype == "application/x-ndjson"
    out = _lines(r)
    assert len(out) == 1 and out[0]["status"] == "error" and out[0]["http_status"] == 500
    assert out[0]["message"].startswith("Failed to load model: ")


@pytest.mark.gpu
@pytest.mark.parametrize("batching", [False, True])
def test_streamed_inference_ends_with_the_plain_answer(cuda, shard_dirs, batching):
    """Partial lines carry the text so far (each a prefix of the next), the last line is the reference's own response
    object and equals what the same request returns without "stream" — one GPU, with and without continuous batching,
    unsharded and as an in-process chain of two shards."""
    case, spec, sd, paths, full = shard_dirs
    env = {"MODEL_CACHE_DIR": os.path.dirname(full)}
    if batching:
        env.update(DLI_BATCHING="1", DLI_MAX_BATCH="2")
    c, w = _client(env)
    prompt = " ".join(str(int(i)) for i in case["prompts"][0])
    T = len(case["prompts"][0])
    reqs = [{"model_name": full, "prompt": prompt, "max_length": T + 40, "do_sample": False}]
    if not batching:
        for i, p in enumerate(paths):
            assert _post(c, "/load_shard", {"model_name": "org/tiny", "shard_id": i, "shard_path": p})[0] == 200
        reqs.append({"model_name": "org/tiny", "prompt": prompt, "shard_ids": [0, 1], "max_length": T + 40, "do_sample": False})
    try:
        for req in reqs:
            code, plain = _post(c, "/inference", req)
            assert code == 200
            out = _lines(c.post("/inference", json=dict(req, stream=True)))
            final = out[-1]
            assert final["http_status"] == 200 and final["status"] == "success" and final["result"] == plain["result"]
            parts = [o["result"] for o in out[:-1]]
            assert parts and all(o["status"] == "partial" for o in out[:-1])
            assert all(b.startswith(a) for a, b in zip(parts, parts[1:] + [final["result"]]))
            assert len(parts[0].split()) > T  # progress, not just the prompt
    finally:
        c.post("/unload_model", json={"model_name": full})
