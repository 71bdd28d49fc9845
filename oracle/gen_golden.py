"""Generate the committed fixtures under tests/golden/ (run in the build container,
where /root/reference and transformers are both present; the fixtures travel, the
reference does not).

  python oracle/gen_golden.py            # writes tests/golden/*.npz, *.json

Three kinds of fixtures:
  1. model cases (*.npz): seeded weights (rebuilt from the seed by
     oracle.decoder_ref.seeded_state_dict, not stored), seeded prompts, and what
     the reference's arithmetic — transformers' forward/generate on those weights —
     produces: bf16 greedy tokens and per-step last-position logits, the same in
     fp32 ("truth"), and hidden states at the 2-shard boundary.
  2. worker_contract.json: request/response/status of the reference's own
     worker/app.py routes, driven in-process with a stub `flask` module
     (Flask is not installed), a stub tokenizer and a tiny model.
  3. shard_ranges.json: layer ranges and metadata.json contents written by the
     reference's own sharder command (shard_model.py) for a grid of (L, N).
"""
from __future__ import annotations

import importlib.util
import json
import os
import sys
import tempfile
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import decoder_ref as R  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
REF = "/root/reference"

CASES = {
    # name: (hf config kwargs, std, batch, prompt_len, new_tokens)
    "llama_tiny": (dict(model_type="llama", vocab_size=1024, hidden_size=512, intermediate_size=1024,
                        num_hidden_layers=4, num_attention_heads=8, num_key_value_heads=2,
                        max_position_embeddings=256, rms_norm_eps=1e-5, rope_theta=10000.0), 0.05, 2, 17, 24),
    "llama_hd128": (dict(model_type="llama", vocab_size=512, hidden_size=512, intermediate_size=768,
                         num_hidden_layers=3, num_attention_heads=4, num_key_value_heads=1,
                         max_position_embeddings=256, rms_norm_eps=1e-5, rope_theta=500000.0), 0.05, 3, 70, 16),
    "mistral_tiny": (dict(model_type="mistral", vocab_size=512, hidden_size=256, intermediate_size=512,
                          num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
                          max_position_embeddings=256, rms_norm_eps=1e-5, rope_theta=10000.0, sliding_window=32),
                     0.06, 1, 48, 16),
    "gpt2_tiny": (dict(model_type="gpt2", vocab_size=1000, n_embd=256, n_layer=2, n_head=4, n_positions=128),
                  0.05, 2, 9, 16),
}


def hf_config(kw):
    from transformers import GPT2Config, LlamaConfig, MistralConfig

    kw = dict(kw)
    mt = kw.pop("model_type")
    cls = {"llama": LlamaConfig, "mistral": MistralConfig, "gpt2": GPT2Config}[mt]
    if mt != "gpt2":
        kw.setdefault("tie_word_embeddings", False)
    return cls(**kw)


def hf_model(cfg, sd, dtype):
    from transformers import AutoModelForCausalLM

    m = AutoModelForCausalLM.from_config(cfg, dtype=dtype).eval()
    sd = {k: v.to(dtype) for k, v in sd.items()}
    if cfg.model_type == "gpt2":
        sd["lm_head.weight"] = sd["transformer.wte.weight"]
    m.load_state_dict(sd, strict=True)
    return m


@torch.no_grad()
def hf_greedy(m, ids, n_new, forced=None):
    """transformers generate(do_sample=False) with per-step logits; with `forced`,
    teacher-forced single-token steps through the same forward + cache."""
    if forced is None:
        out = m.generate(ids, attention_mask=torch.ones_like(ids), max_length=ids.shape[1] + n_new, do_sample=False,
                         pad_token_id=0, output_logits=True, return_dict_in_generate=True)
        return out.sequences[:, ids.shape[1]:], torch.stack([l.float() for l in out.logits])
    logits, past, cur = [], None, ids
    for i in range(n_new):
        o = m(cur, past_key_values=past, use_cache=True)
        past = o.past_key_values
        logits.append(o.logits[:, -1].float())
        cur = forced[:, i:i + 1]
    return forced, torch.stack(logits)


def gen_model_cases():
    for name, (kw, std, B, T, N) in CASES.items():
        cfg = hf_config(kw)
        spec = R.ModelSpec.from_hf(cfg)
        seed = 11
        sd = R.seeded_state_dict(spec, seed=seed, std=std, dtype=torch.bfloat16)
        ids = R.seeded_prompts(spec.vocab, B, T, seed=seed + 1)
        m16 = hf_model(cfg, sd, torch.bfloat16)
        m32 = hf_model(cfg, sd, torch.float32)
        tok16, lg16 = hf_greedy(m16, ids, N)
        _, lg32 = hf_greedy(m32, ids, N, forced=tok16)  # fp32 truth along the bf16 path
        with torch.no_grad():
            hs = m16(ids, output_hidden_states=True).hidden_states
        mid = R.shard_ranges(spec.n_layers, 2)[0][1] + 1  # hidden_states[k] = input of layer k
        top2 = lg32.topk(2, dim=-1).values
        margin32 = (top2[..., 0] - top2[..., 1])
        err16 = (lg16 - lg32).abs().amax(dim=-1)
        np.savez_compressed(
            os.path.join(GOLD, f"{name}.npz"),
            config=json.dumps(kw), seed=seed, std=std, prompts=ids.numpy(),
            hf_bf16_tokens=tok16.numpy(),
            hf_bf16_logits=lg16.to(torch.bfloat16).view(torch.int16).numpy(),  # [N,B,V] raw bf16 bits
            hf_fp32_logits=lg32.to(torch.float16).numpy(),
            hf_fp32_margin=margin32.numpy(), hf_bf16_err=err16.numpy(),
            boundary_hidden=hs[mid].to(torch.bfloat16).view(torch.int16).numpy(), boundary_layer=mid,
        )
        agree = (lg16.argmax(-1) == lg32.argmax(-1)).float().mean().item()
        print(f"{name}: min fp32 margin {margin32.min():.4f}, max bf16 err {err16.max():.4f}, "
              f"argmax bf16==fp32 {agree:.3f}")


# ------------------------------------------------------------------ reference worker contract
def flask_stub():
    """Minimal in-memory `flask` so the unmodified reference worker/app.py imports."""
    mod = types.ModuleType("flask")

    class _Req:
        json = None
        headers = {}

    req = _Req()

    class Flask:
        def __init__(self, *_a, **_k):
            self.routes = {}

        def route(self, path, methods=("GET",)):
            def deco(f):
                self.routes[path] = f
                return f
            return deco

        def run(self, *a, **k):
            pass

    mod.Flask, mod.request = Flask, req
    mod.jsonify = lambda obj: obj
    return mod


class StubTokenizer:
    """Whitespace-separated integer ids <-> text (no tokenizer files offline)."""

    def __call__(self, prompt, return_tensors="pt"):
        ids = torch.tensor([[int(t) for t in prompt.split()]], dtype=torch.int64)
        return {"input_ids": ids, "attention_mask": torch.ones_like(ids)}

    def decode(self, ids, skip_special_tokens=True):
        return " ".join(str(int(i)) for i in ids)


def load_reference_worker(env=None):
    os.environ.update(env or {})
    os.environ.setdefault("MODEL_CACHE_DIR", tempfile.mkdtemp(prefix="dli_ref_cache_"))
    sys.modules["flask"] = flask_stub()
    spec = importlib.util.spec_from_file_location("ref_worker_app", os.path.join(REF, "worker", "app.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def call(mod, path, body=None, headers=None):
    import flask

    flask.request.json = body
    flask.request.headers = headers or {}
    r = mod.app.routes[path]()
    status = 200
    if isinstance(r, tuple):
        r, status = r
    return status, r


def gen_worker_contract():
    from transformers import AutoModelForCausalLM

    mod = load_reference_worker({"USE_GPU": "0", "AUTH_ENABLED": "0"})
    cfg = hf_config(CASES["gpt2_tiny"][0])
    torch.manual_seed(0)
    model = AutoModelForCausalLM.from_config(cfg).eval()
    tmp = tempfile.mkdtemp(prefix="dli_ref_shard_")
    shard_dir = os.path.join(tmp, "tiny", "shard_1")
    model.save_pretrained(shard_dir)
    meta = {"model_name": "tiny", "shard_id": 1, "num_shards": 2, "start_layer": 1, "end_layer": 1, "total_layers": 2}
    json.dump(meta, open(os.path.join(shard_dir, "metadata.json"), "w"))
    mod.loaded_tokenizers["tiny"] = StubTokenizer()  # the hub is offline: pre-seed the tokenizer dict
    rec = []

    def do(path, body=None, note=""):
        st, resp = call(mod, path, body)
        if isinstance(resp, dict):
            resp = json.loads(json.dumps(resp, default=str))
            if "execution_time" in resp:
                resp["execution_time"] = "<float>"
            if "resources" in resp:
                resp["resources"] = {k: ("<float>" if isinstance(v, float) else v) for k, v in resp["resources"].items()}
            if path == "/inference" and st == 200:
                resp["result"] = "<str>"
        rec.append({"route": path, "request": body, "status": st, "response": resp, "note": note})

    do("/health", None, "empty worker")
    do("/load_model", {}, "missing model_name")
    do("/load_shard", {"model_name": "tiny", "shard_id": 0, "shard_path": shard_dir}, "shard_id 0 is falsy (Q1)")
    do("/load_shard", {"model_name": "tiny", "shard_id": 1, "shard_path": shard_dir})
    do("/load_shard", {"model_name": "tiny", "shard_id": 1, "shard_path": shard_dir}, "idempotent")
    do("/load_shard", {"model_name": "tiny", "shard_id": 1}, "missing shard_path")
    do("/health", None, "one shard registered")
    do("/inference", {"model_name": "tiny"}, "missing prompt")
    do("/inference", {"model_name": "tiny", "prompt": "1 2 3", "shard_ids": [2]}, "shard not loaded")
    do("/inference", {"model_name": "tiny", "prompt": "1 2 3 4", "shard_ids": [1], "max_length": 12})
    do("/inference", {"model_name": "tiny", "prompt": "1 2 3 4", "shard_ids": [1], "max_length": 12, "timeout": 0},
       "sharded path never checks the timeout")
    mod.loaded_models["tiny"] = model
    do("/inference", {"model_name": "tiny", "prompt": "1 2 3 4", "max_length": 12}, "unsharded path")
    do("/inference", {"model_name": "tiny", "prompt": "1 2 3 4", "max_length": 12, "timeout": -1}, "timeout -> 408")
    do("/unload_model", {}, "missing model_name")
    do("/unload_model", {"model_name": "tiny"})
    do("/unload_model", {"model_name": "never-loaded"}, "unloading nothing succeeds")
    do("/health", None, "after unload: lazily loaded tiny_1 leaks (Q7)")
    # auth
    mod.AUTH_ENABLED, mod.AUTH_KEY = True, "k"
    st, resp = call(mod, "/health", None, {})
    rec.append({"route": "/health", "request": None, "status": st, "response": resp, "note": "AUTH_ENABLED=1, no header"})
    st, resp = call(mod, "/health", None, {"Authorization": "Bearer k"})
    rec.append({"route": "/health", "request": None, "status": st, "response": {"status": resp["status"]},
                "note": "AUTH_ENABLED=1, Bearer k"})
    json.dump(rec, open(os.path.join(GOLD, "worker_contract.json"), "w"), indent=1)
    print(f"worker_contract.json: {len(rec)} exchanges")


# ------------------------------------------------------------------ reference sharder
def gen_shard_ranges():
    """Run the reference's shard_model command on tiny GPT-2 models (its only working
    branch) and record the metadata.json files it writes."""
    import transformers

    dj = types.ModuleType("django")
    for name in ("django.core", "django.core.management", "django.core.management.base"):
        sys.modules[name] = types.ModuleType(name)
    sys.modules["django"] = dj

    class BaseCommand:
        class _Style:
            SUCCESS = ERROR = staticmethod(lambda s: s)

        style = _Style()

        class _Out:
            @staticmethod
            def write(_s):
                pass

        stdout = _Out()

    sys.modules["django.core.management.base"].BaseCommand = BaseCommand
    path = os.path.join(REF, "master", "dashboard", "management", "commands", "shard_model.py")
    spec = importlib.util.spec_from_file_location("ref_shard_model", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)

    class _Tok:
        @staticmethod
        def from_pretrained(*_a, **_k):
            class T:
                def save_pretrained(self, p):
                    open(os.path.join(p, "stub.txt"), "w").write("stub")
            return T()

    out = []
    for L, N in [(12, 1), (12, 2), (12, 4), (12, 5), (6, 4), (22, 2), (32, 4), (32, 3), (80, 8), (7, 7), (5, 3)]:
        cfg = transformers.GPT2Config(vocab_size=64, n_embd=32, n_layer=L, n_head=2, n_positions=16)

        class _Model:
            @staticmethod
            def from_pretrained(*_a, **_k):
                return transformers.GPT2LMHeadModel(cfg)

        mod.AutoTokenizer, mod.AutoModelForCausalLM = _Tok, _Model
        tmp = tempfile.mkdtemp(prefix="dli_ref_sharder_")
        mod.Command().handle(model_name="org/tiny", num_shards=N, output_dir=tmp)
        metas = [json.load(open(os.path.join(tmp, "org_tiny", f"shard_{i}", "metadata.json"))) for i in range(N)]
        out.append({"n_layers": L, "n_shards": N, "dir": "org_tiny", "has_tokenizer_dir":
                    os.path.isdir(os.path.join(tmp, "org_tiny", "tokenizer")), "metadata": metas})
    json.dump(out, open(os.path.join(GOLD, "shard_ranges.json"), "w"), indent=1)
    print(f"shard_ranges.json: {len(out)} (L, N) cases")


if __name__ == "__main__":
    os.makedirs(GOLD, exist_ok=True)
    os.environ.setdefault("TRANSFORMERS_OFFLINE", "1")
    which = sys.argv[1:] or ["models", "worker", "shards"]
    if "models" in which:
        gen_model_cases()
    if "worker" in which:
        gen_worker_contract()
    if "shards" in which:
        gen_shard_ranges()
