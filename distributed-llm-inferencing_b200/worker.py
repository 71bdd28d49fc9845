"""Drop-in for the reference worker service (reference worker/app.py).

Same routes, methods, JSON keys, status codes and environment variables as the
Flask app the master talks to (worker/app.py:49-330; contract table SURVEY.md
§A.1; behaviour recorded from the reference itself in
tests/golden/worker_contract.json):

    GET  /health          worker/app.py:49-92
    POST /load_model      worker/app.py:94-137
    POST /load_shard      worker/app.py:139-206
    POST /unload_model    worker/app.py:208-250
    POST /inference       worker/app.py:252-330  (+ run_shard_inference :332-372)

`/ssh_setup` (:374-413, a paramiko login test) is out of scope.  Behind the routes
the model object is a ShardEngine on the B200 instead of a transformers model on
the CPU.  Built on werkzeug (Flask is not installed; werkzeug is what Flask wraps).

Deliberate deviations from the reference, each a fixed defect:
  Q1  `shard_id: 0` is accepted (the reference's `all([...])` rejects falsy ids);
  Q2  every shard id of the request is used (the reference silently uses only the
      first), as an in-process layer-range chain; the shards must cover all layers;
  Q3  sampling like the reference by default (do_sample=True, top_k=50, top_p=0.95,
      temperature=0.8) but reproducible: optional request fields `seed` and
      `do_sample` (false = greedy, the parity mode);
  Q6  the timeout is checked between decode steps, not only around generate();
  Q7  /unload_model also frees the lazily created per-shard engines;
  USE_GPU defaults to 1: this worker has no CPU path (the reference defaults to 0).
"""
from __future__ import annotations

import json
import os
import threading
import time
from typing import Dict

import psutil
import torch
from werkzeug.wrappers import Request, Response

from .sharding import load_hf_state_dict, read_metadata


class IdTokenizer:
    """Used when no tokenizer files exist (the hub is unreachable): prompts are
    whitespace-separated token ids."""

    def __call__(self, prompt, return_tensors="pt"):
        ids = torch.tensor([[int(t) for t in str(prompt).split()]], dtype=torch.int64)
        return {"input_ids": ids, "attention_mask": torch.ones_like(ids)}

    def decode(self, ids, skip_special_tokens=True):
        return " ".join(str(int(i)) for i in ids)


def _load_tokenizer(path: str):
    names = ("tokenizer.json", "tokenizer.model", "vocab.json", "tokenizer_config.json")
    if path and os.path.isdir(path) and any(os.path.exists(os.path.join(path, n)) for n in names):
        from transformers import AutoTokenizer

        return AutoTokenizer.from_pretrained(path)
    return IdTokenizer()


def _stop_tokens(path: str, cfg: dict):
    """(eos ids, pad id) generate() stops on: generation_config.json wins over config.json, as in
    transformers (the reference calls model.generate with the checkpoint's generation config)."""
    gen = {}
    gp = os.path.join(path, "generation_config.json")
    if os.path.exists(gp):
        with open(gp) as f:
            gen = json.load(f)
    eos = gen.get("eos_token_id", cfg.get("eos_token_id"))
    pad = gen.get("pad_token_id", cfg.get("pad_token_id"))
    if eos is None:
        eos = ()
    elif isinstance(eos, int):
        eos = (eos,)
    return tuple(int(e) for e in eos), (None if pad is None else int(pad))


class Worker:
    """State + handlers; one instance per process (the reference keeps the same
    three dicts as module globals, worker/app.py:14-16)."""

    def __init__(self, env=None):
        env = os.environ if env is None else env
        self.loaded_models: Dict[str, object] = {}
        self.loaded_tokenizers: Dict[str, object] = {}
        self.loaded_shards: Dict[str, Dict[object, dict]] = {}
        self.model_cache_dir = env.get("MODEL_CACHE_DIR", "/app/model_cache")
        self.use_gpu = env.get("USE_GPU", "1") == "1"
        self.device = "cuda" if torch.cuda.is_available() and self.use_gpu else "cpu"
        self.cuda_device = "cuda:0"  # a worker group binds every rank to its own GPU (pipeline_worker.py)
        self.auth_enabled = env.get("AUTH_ENABLED", "0") == "1"
        self.auth_key = env.get("AUTH_KEY", "")
        self.max_seq_len = int(env.get("DLI_MAX_SEQ_LEN", "4096"))
        # prompts longer than this many tokens are prefilled in chunks (0 = whole prompt at once)
        self.prefill_chunk = int(env.get("DLI_PREFILL_CHUNK", "0")) or None
        # DLI_BATCHING=1: concurrent /inference requests on an unsharded model share decode steps
        # (continuous batching; greedy or the reference's sampling call per request); the reference serialises requests (worker/Dockerfile:45)
        self.batching = env.get("DLI_BATCHING", "0") == "1"
        self.max_batch = int(env.get("DLI_MAX_BATCH", "8")) if self.batching else 1
        # DLI_KV_POOL_TOKENS: size of the paged KV pool in tokens (default: max batch x max sequence length, i.e. every
        # sequence can reach the maximum at once); with batching, sequences take pages from it as they grow
        self.kv_pool_tokens = int(env.get("DLI_KV_POOL_TOKENS", "0")) or None
        self.batchers: Dict[str, object] = {}
        # With batching the WSGI server is threaded: everything that touches the dicts above or runs an
        # engine directly holds this lock; only the wait on a batcher's future happens outside it.
        self._lock = threading.RLock()
        self._progress = threading.local()  # .cb(text) while the calling thread serves a streaming /inference
        self.routes = {("GET", "/health"): self.health, ("POST", "/load_model"): self.load_model,
                       ("POST", "/load_shard"): self.load_shard, ("POST", "/unload_model"): self.unload_model,
                       ("POST", "/inference"): self.inference}

    # ------------------------------------------------------------------ helpers
    def _resolve_model_dir(self, model_name: str) -> str:
        for cand in (model_name, os.path.join(self.model_cache_dir, model_name.replace("/", "_")),
                     os.path.join(self.model_cache_dir, model_name)):
            if os.path.isdir(cand) and os.path.exists(os.path.join(cand, "config.json")):
                return cand
        raise FileNotFoundError(f"{model_name} is not a local model directory and the Hugging Face hub is not "
                                f"reachable (looked in {self.model_cache_dir})")

    def _engine(self, path: str, start=0, end=None, max_batch=None, head_shard=None, head_path=None):
        """head_shard=(rank, world) + head_path: a stage of a worker group with the lm_head split by vocabulary rows
        over the stages; the final norm and this stage's rows of the head are read from `head_path`, the directory of
        the LAST shard (the only one that stores them)."""
        from .engine import ModelSpec, ShardEngine
        from .sharding import load_named_tensors

        if self.device != "cuda":
            raise RuntimeError("this worker has no CPU path: it needs a CUDA device and USE_GPU=1")
        cfg = json.load(open(os.path.join(path, "config.json")))
        spec = ModelSpec.from_hf_config(cfg)
        if max_batch is None:
            max_batch = self.max_batch if end is None else 1
        eng = ShardEngine(spec, start, end, device=self.cuda_device, max_batch=max_batch,
                          max_seq_len=min(self.max_seq_len, spec.max_pos), kv_pool_tokens=self.kv_pool_tokens,
                          head_shard=head_shard)
        sd = load_hf_state_dict(path, eng.start, eng.end, spec.n_layers)
        if head_shard is not None:
            sd.update(load_named_tensors(head_path, ("model.norm.weight", "lm_head.weight") + (
                ("model.embed_tokens.weight",) if spec.tie_embeddings else ())))
        eng.load_state_dict(sd)
        eng.eos_token_ids, eng.pad_token_id = _stop_tokens(path, cfg)
        eng.prefill_chunk = self.prefill_chunk
        return eng

    # ------------------------------------------------------------------ routes
    def health(self, _data):
        gpu = 0.0
        if torch.cuda.is_available() and self.use_gpu:
            mx = torch.cuda.max_memory_allocated()
            gpu = torch.cuda.memory_allocated() / mx if mx > 0 else 0.0
        shard_info = [{"model_name": m, "shard_id": sid, "path": d["path"], "metadata": d["metadata"]}
                      for m, shards in self.loaded_shards.items() for sid, d in shards.items()]
        return {"status": "healthy",
                "resources": {"cpu": psutil.cpu_percent() / 100.0, "memory": psutil.virtual_memory().percent / 100.0,
                              "gpu": gpu, "gpu_available": torch.cuda.is_available() and self.use_gpu,
                              "device": self.device},
                "loaded_models": list(self.loaded_models.keys()),
                "loaded_tokenizers": list(self.loaded_tokenizers.keys()),
                "loaded_shards": shard_info}, 200

    def load_model(self, data):
        model_name = (data or {}).get("model_name")
        if not model_name:
            return {"status": "error", "message": "Model name is required"}, 400
        try:
            if model_name in self.loaded_models:
                return {"status": "success", "message": f"Model {model_name} is already loaded"}, 200
            path = self._resolve_model_dir(model_name)
            if model_name not in self.loaded_tokenizers:
                self.loaded_tokenizers[model_name] = _load_tokenizer(path)
            self.loaded_models[model_name] = self._engine(path)
            if self.batching:
                from .batching import ContinuousBatcher

                self.batchers[model_name] = ContinuousBatcher(self.loaded_models[model_name]).start()
            return {"status": "success", "message": f"Model {model_name} loaded successfully on {self.device}"}, 200
        except Exception as e:  # noqa: BLE001 — the reference maps every failure to 500
            return {"status": "error", "message": f"Failed to load model: {str(e)}"}, 500

    def load_shard(self, data):
        data = data or {}
        model_name, shard_id, shard_path = data.get("model_name"), data.get("shard_id"), data.get("shard_path")
        if not model_name or shard_id is None or shard_id == "" or not shard_path:  # Q1: 0 is a valid id
            return {"status": "error", "message": "Model name, shard ID, and shard path are required"}, 400
        try:
            shards = self.loaded_shards.setdefault(model_name, {})
            if shard_id in shards:
                return {"status": "success", "message": f"Shard {shard_id} of model {model_name} is already loaded"}, 200
            if model_name not in self.loaded_tokenizers:
                self.loaded_tokenizers[model_name] = _load_tokenizer(os.path.join(os.path.dirname(shard_path), "tokenizer"))
            shards[shard_id] = {"path": shard_path, "metadata": read_metadata(shard_path, model_name, shard_id)}
            return {"status": "success", "message": f"Shard {shard_id} of model {model_name} loaded successfully"}, 200
        except Exception as e:  # noqa: BLE001
            return {"status": "error", "message": f"Failed to load shard: {str(e)}"}, 500

    def unload_model(self, data):
        model_name = (data or {}).get("model_name")
        if not model_name:
            return {"status": "error", "message": "Model name is required"}, 400
        try:
            if model_name in self.batchers:
                self.batchers.pop(model_name).stop()
            # Q7: the lazily created per-shard engines too — keyed by the shard ids registered for THIS model, so
            # another model whose name merely starts with "<model_name>_" stays loaded (the reference deletes the
            # exact key only, worker/app.py:223-227)
            keys = {model_name} | {f"{model_name}_{sid}" for sid in self.loaded_shards.get(model_name, {})}
            for key in [k for k in self.loaded_models if k in keys]:
                self._release_engine(key, self.loaded_models.pop(key))
            self.loaded_tokenizers.pop(model_name, None)
            self.loaded_shards.pop(model_name, None)
            if self.device == "cuda":
                torch.cuda.empty_cache()
            return {"status": "success", "message": f"Model {model_name} unloaded successfully"}, 200
        except Exception as e:  # noqa: BLE001
            return {"status": "error", "message": f"Failed to unload model: {str(e)}"}, 500

    def _on_tokens(self, tokenizer):
        """Progress callback for the generate paths of the calling thread, or None when nobody is streaming."""
        cb = getattr(self._progress, "cb", None)
        if cb is None:
            return None
        return lambda ids: cb(tokenizer.decode(ids[0], skip_special_tokens=True))

    def _wait_streaming(self, fut, tokenizer, prompt_ids, new_tokens):
        """Wait for a batcher / worker-group future; while streaming, report the tokens it has so far."""
        from concurrent.futures import wait

        on = self._on_tokens(tokenizer)
        while on is not None and not fut.done():
            wait([fut], timeout=0.05)
            toks = list(new_tokens())
            if toks:
                on(torch.cat((prompt_ids.long(), torch.tensor(toks, dtype=torch.int64)))[None])
        return fut

    def _release_engine(self, key, eng) -> None:
        """Free an engine that /unload_model has just removed (weights, KV pages, captured graphs)."""
        if hasattr(eng, "unload"):
            eng.unload()

    def inference(self, data):
        data = data or {}
        model_name, prompt = data.get("model_name"), data.get("prompt")
        max_length = data.get("max_length", 100)
        shard_ids = data.get("shard_ids")
        timeout = data.get("timeout", 60)
        if not model_name or not prompt:
            return {"status": "error", "message": "Model name and prompt are required"}, 400
        try:
            start_time = time.time()
            deadline = start_time + timeout
            if shard_ids:
                with self._lock:
                    result = self.run_shard_inference(model_name, prompt, shard_ids, max_length, deadline,
                                                      self._gen_kwargs(data))
            else:
                with self._lock:
                    if model_name not in self.loaded_models:
                        resp, code = self.load_model({"model_name": model_name})
                        if code == 500:
                            return resp, code
                    model, tokenizer = self.loaded_models[model_name], self.loaded_tokenizers[model_name]
                    batcher = self.batchers.get(model_name)
                    ids = tokenizer(prompt, return_tensors="pt")["input_ids"]
                    if (time.time() - start_time) > timeout:
                        raise TimeoutError("Inference preparation timed out")
                    if batcher is None:
                        outputs = model.generate(ids, max_length=max_length, deadline=deadline,
                                                 on_tokens=self._on_tokens(tokenizer), **self._gen_kwargs(data))
                if batcher is not None:  # joins the running batch; the wait is outside the lock
                    gen = self._gen_kwargs(data)
                    fut = batcher.submit(ids[0], max_length, timeout, do_sample=gen["do_sample"], seed=gen["seed"])
                    self._wait_streaming(fut, tokenizer, ids[0], lambda: getattr(fut, "request").out)
                    outputs = fut.result()[None]
                result = tokenizer.decode(outputs[0], skip_special_tokens=True)
                if (time.time() - start_time) > timeout:
                    raise TimeoutError("Inference generation timed out")
            return {"status": "success", "result": result, "execution_time": time.time() - start_time}, 200
        except TimeoutError as e:
            return {"status": "error", "message": str(e)}, 408
        except Exception as e:  # noqa: BLE001
            return {"status": "error", "message": f"Inference failed: {str(e)}"}, 500

    @staticmethod
    def _gen_kwargs(data):
        """generate() arguments of worker/app.py:297-305, plus the optional overrides."""
        return dict(do_sample=bool(data.get("do_sample", True)), temperature=0.8, top_k=50, top_p=0.95,
                    seed=int(data.get("seed", int(time.time() * 1e6) & 0x7fffffff)))

    def run_shard_inference(self, model_name, prompt, shard_ids, max_length, deadline=None, gen=None):
        """worker/app.py:332-372, with the intended semantics: the listed shards are
        layer ranges run in order on hidden states (lazy weight load on first use)."""
        from .engine import LocalChain

        engines = []
        for shard_id in shard_ids:
            if shard_id not in self.loaded_shards.get(model_name, {}):
                raise ValueError(f"Shard {shard_id} of model {model_name} is not loaded")
            info = self.loaded_shards[model_name][shard_id]
            key = f"{model_name}_{shard_id}"
            if key not in self.loaded_models:
                md = info["metadata"]
                self.loaded_models[key] = self._engine(info["path"], md["start_layer"], md["end_layer"])
            engines.append(self.loaded_models[key])
        tokenizer = self.loaded_tokenizers[model_name]
        ids = tokenizer(prompt, return_tensors="pt")["input_ids"]
        outputs = LocalChain(engines).generate(ids, max_length=max_length, deadline=deadline,
                                               on_tokens=self._on_tokens(tokenizer), **(gen or {}))
        return tokenizer.decode(outputs[0], skip_special_tokens=True)

    # ------------------------------------------------------------------ WSGI
    def dispatch(self, method: str, path: str, data, headers):
        if self.auth_enabled:  # worker/app.py:32-47
            if headers.get("Authorization") != f"Bearer {self.auth_key}":
                return {"status": "error", "message": "Unauthorized access"}, 401
        fn = self.routes.get((method, path))
        if fn is None:
            known = {p for _, p in self.routes}
            return {"status": "error", "message": "Method not allowed" if path in known else "Not found"}, \
                (405 if path in known else 404)
        if fn == self.inference:  # takes the lock itself, around everything but a batcher wait
            return fn(data)
        with self._lock:
            return fn(data)

    def stream_inference(self, data, headers):
        """`"stream": true` on /inference (an extension; the reference answers with one JSON body, worker/app.py:314-318):
        newline-delimited JSON — {"status": "partial", "result": <text so far>} as tokens arrive (every 8 decode steps
        on one GPU, every block of rounds in a worker group), then the reference's own final object (or error object)
        with its HTTP status under "http_status"."""
        import queue

        q: "queue.Queue" = queue.Queue()
        plain = {k: v for k, v in data.items() if k != "stream"}
        hdrs = dict(headers)

        def work():
            self._progress.cb = lambda text: q.put(("partial", text))
            try:
                out = self.dispatch("POST", "/inference", plain, hdrs)
            except Exception as e:  # noqa: BLE001
                out = ({"status": "error", "message": f"Inference failed: {e}"}, 500)
            finally:
                self._progress.cb = None
            q.put(("final", out))

        threading.Thread(target=work, daemon=True).start()

        def lines():
            last = None
            while True:
                kind, v = q.get()
                if kind == "partial":
                    if v != last:
                        last = v
                        yield json.dumps({"status": "partial", "result": v}) + "\n"
                else:
                    body, code = v
                    yield json.dumps(dict(body, http_status=code)) + "\n"
                    return

        return Response(lines(), status=200, mimetype="application/x-ndjson")

    def __call__(self, environ, start_response):
        req = Request(environ)
        data = req.get_json(silent=True) if req.method == "POST" else None
        if req.method == "POST" and req.path == "/inference" and isinstance(data, dict) and data.get("stream"):
            if self.auth_enabled and req.headers.get("Authorization") != f"Bearer {self.auth_key}":
                body, code = {"status": "error", "message": "Unauthorized access"}, 401
            elif not data.get("model_name") or not data.get("prompt"):
                body, code = {"status": "error", "message": "Model name and prompt are required"}, 400
            else:
                return self.stream_inference(data, req.headers)(environ, start_response)
            return Response(json.dumps(body), status=code, mimetype="application/json")(environ, start_response)
        body, code = self.dispatch(req.method, req.path, data, req.headers)
        return Response(json.dumps(body), status=code, mimetype="application/json")(environ, start_response)


def main(argv=None):
    import argparse

    from werkzeug.serving import run_simple

    ap = argparse.ArgumentParser(description="B200 drop-in for the reference worker (port 5000 like worker/app.py:416)")
    ap.add_argument("--host", default="0.0.0.0")
    ap.add_argument("--port", type=int, default=5000)
    a = ap.parse_args(argv)
    # one request at a time, like the reference's single sync gunicorn worker (worker/Dockerfile:45)
    w = Worker()
    run_simple(a.host, a.port, w, threaded=w.batching)


if __name__ == "__main__":
    main()
