"""Shard assignment and shard directories.

Keeps the reference's shard-assignment semantics
(master/dashboard/management/commands/shard_model.py:55-66, :94-109):
  * `L // N` layers per shard, the whole remainder on the LAST shard;
  * directory layout `<out>/<model_name with '/' -> '_'>/shard_<i>/` with a
    sibling `tokenizer/` directory (consumed at worker/app.py:168,:177);
  * `metadata.json` = {model_name, shard_id, num_shards, start_layer,
    end_layer (inclusive), total_layers}.
What it does NOT keep is the reference's storage defect: the
reference saves a full-depth model per shard with random out-of-range layers.
`write_shards` stores only the tensors of the shard's layer range (+ embedding on
shard 0, final norm / lm_head on the last shard), and `load_hf_state_dict` reads
only in-range tensors from ANY Hugging Face directory — including the reference's
full-depth shard directories, whose out-of-range tensors are never trusted.
"""
from __future__ import annotations

import json
import os
import re
from typing import Dict, List, Optional, Tuple

import torch

_LAYER_RE = re.compile(r"^(?:model\.layers|transformer\.h)\.(\d+)\.")
_FIRST_ONLY = ("model.embed_tokens.", "transformer.wte.", "transformer.wpe.")
_LAST_ONLY = ("model.norm.", "lm_head.", "transformer.ln_f.")


def shard_ranges(n_layers: int, n_shards: int) -> List[Tuple[int, int]]:
    """Inclusive (start_layer, end_layer) of every shard (shard_model.py:55-66)."""
    if n_shards < 1 or n_shards > n_layers:
        raise ValueError(f"cannot split {n_layers} layers into {n_shards} shards")
    per, rem = n_layers // n_shards, n_layers % n_shards
    out = []
    for i in range(n_shards):
        start = i * per
        end = start + per + (rem if i == n_shards - 1 else 0)
        out.append((start, end - 1))
    return out


def shard_metadata(model_name: str, shard_id: int, n_shards: int, n_layers: int) -> Dict[str, object]:
    """The metadata.json the reference writes (shard_model.py:99-106)."""
    start, end = shard_ranges(n_layers, n_shards)[shard_id]
    return {"model_name": model_name, "shard_id": shard_id, "num_shards": n_shards, "start_layer": start,
            "end_layer": end, "total_layers": n_layers}


def default_metadata(model_name: str, shard_id) -> Dict[str, object]:
    """What the reference worker synthesises when metadata.json is absent (worker/app.py:181-189)."""
    return {"model_name": model_name, "shard_id": shard_id, "num_shards": 1, "start_layer": 0, "end_layer": 0,
            "total_layers": 1}


def model_dir_name(model_name: str) -> str:
    return model_name.replace("/", "_")  # shard_model.py:25


def tensor_in_range(name: str, start: int, end: int, is_first: bool, is_last: bool, tied: bool = False) -> bool:
    """Does a HF tensor name belong to the stage owning layers [start, end]?"""
    m = _LAYER_RE.match(name)
    if m:
        return start <= int(m.group(1)) <= end
    if name.startswith(_FIRST_ONLY):
        # a tied lm_head reads the embedding on the last stage too (GPT-2, tie_word_embeddings)
        return is_first or (is_last and (tied or name.startswith("transformer.wte.")))
    if name.startswith(_LAST_ONLY):
        return is_last
    return False


def _tensor_files(path: str) -> List[str]:
    for index in ("model.safetensors.index.json", "pytorch_model.bin.index.json"):
        p = os.path.join(path, index)
        if os.path.exists(p):
            files = sorted(set(json.load(open(p))["weight_map"].values()))
            return [os.path.join(path, f) for f in files]
    for f in ("model.safetensors", "pytorch_model.bin"):
        p = os.path.join(path, f)
        if os.path.exists(p):
            return [p]
    raise FileNotFoundError(f"no model.safetensors / pytorch_model.bin (or index) in {path}")


def load_hf_state_dict(path: str, start_layer: int = 0, end_layer: Optional[int] = None,
                       total_layers: Optional[int] = None) -> Dict[str, torch.Tensor]:
    """Read only the tensors a stage needs from a Hugging Face model directory
    (`save_pretrained` format: safetensors or .bin, single file or sharded index)."""
    cfg = json.load(open(os.path.join(path, "config.json")))
    L = total_layers or cfg.get("num_hidden_layers") or cfg.get("n_layer")
    end = L - 1 if end_layer is None else end_layer
    is_first, is_last = start_layer == 0, end == L - 1
    tied = bool(cfg.get("tie_word_embeddings", cfg.get("model_type") == "gpt2"))
    out: Dict[str, torch.Tensor] = {}
    for f in _tensor_files(path):
        if f.endswith(".safetensors"):
            from safetensors import safe_open

            with safe_open(f, framework="pt", device="cpu") as sf:
                for name in sf.keys():
                    if tensor_in_range(name, start_layer, end, is_first, is_last, tied):
                        out[name] = sf.get_tensor(name)
        else:
            sd = torch.load(f, map_location="cpu", weights_only=True)
            for name, t in sd.items():
                if tensor_in_range(name, start_layer, end, is_first, is_last, tied):
                    out[name] = t
    return out


def load_named_tensors(path: str, names) -> Dict[str, torch.Tensor]:
    """Only the tensors called `names` from a Hugging Face directory (those that exist there)."""
    want, out = set(names), {}
    for f in _tensor_files(path):
        if f.endswith(".safetensors"):
            from safetensors import safe_open

            with safe_open(f, framework="pt", device="cpu") as sf:
                for name in sf.keys():
                    if name in want:
                        out[name] = sf.get_tensor(name)
        else:
            sd = torch.load(f, map_location="cpu", weights_only=True)
            out.update({k: v for k, v in sd.items() if k in want})
    return out


def write_shards(state_dict: Dict[str, torch.Tensor], config: Dict[str, object], model_name: str, n_shards: int,
                 output_dir: str = "model_shards") -> List[str]:
    """Real shard materialiser: the same CLI contract as
    `manage.py shard_model --model_name --num_shards --output_dir`
    (shard_model.py:11-14), but each shard directory holds only its own tensors."""
    from safetensors.torch import save_file

    L = config.get("num_hidden_layers") or config.get("n_layer")
    tied = bool(config.get("tie_word_embeddings", config.get("model_type") == "gpt2"))
    base = os.path.join(output_dir, model_dir_name(model_name))
    os.makedirs(os.path.join(base, "tokenizer"), exist_ok=True)
    paths = []
    for i, (a, b) in enumerate(shard_ranges(L, n_shards)):
        p = os.path.join(base, f"shard_{i}")
        os.makedirs(p, exist_ok=True)
        part = {k: v.contiguous() for k, v in state_dict.items()
                if tensor_in_range(k, a, b, a == 0, b == L - 1, tied)}
        save_file(part, os.path.join(p, "model.safetensors"))
        json.dump(config, open(os.path.join(p, "config.json"), "w"), indent=2)
        json.dump(shard_metadata(model_name, i, n_shards, L), open(os.path.join(p, "metadata.json"), "w"), indent=2)
        paths.append(p)
    return paths


_TOKENIZER_FILES = ("tokenizer.json", "tokenizer.model", "tokenizer_config.json", "special_tokens_map.json",
                    "vocab.json", "merges.txt", "added_tokens.json")


def _iter_source_tensors(path: str):
    """(name, loader) for every tensor of a Hugging Face directory, one file at a time; safetensors tensors are
    read individually on demand, so the caller never holds more than it asks for."""
    for f in _tensor_files(path):
        if f.endswith(".safetensors"):
            from safetensors import safe_open

            with safe_open(f, framework="pt", device="cpu") as sf:
                for name in sf.keys():
                    yield name, (lambda sf=sf, name=name: sf.get_tensor(name))
        else:  # a .bin checkpoint can only be read whole, file by file
            sd = torch.load(f, map_location="cpu", weights_only=True)
            for name in list(sd):
                yield name, (lambda sd=sd, name=name: sd[name])


def shard_model_dir(source_dir: str, model_name: str, n_shards: int, output_dir: str = "model_shards",
                    max_file_bytes: int = 2 << 30) -> List[str]:
    """The reference's `manage.py shard_model` (shard_model.py:16-114) as a streaming tool: split the model in
    `source_dir` (a `save_pretrained` directory: safetensors or .bin, single file or sharded index) into
    `n_shards` layer-range shard directories under `<output_dir>/<model_name with / -> _>/shard_<i>/`, each with
    config.json, the reference's metadata.json and ONLY its own tensors (+ embedding on shard 0, final norm / lm_head
    on the last; a tied head's embedding on both), plus the sibling `tokenizer/` directory.  Tensors are copied one
    at a time into output files of at most `max_file_bytes` (with a `model.safetensors.index.json`), so a 70B
    checkpoint (141 GB) is sharded with a few GB of RAM — the reference instantiates the full model once per shard
    (shard_model.py:37,71).  Works for every architecture whose layers are `model.layers.N.` or `transformer.h.N.`
    (Llama, Mistral, TinyLlama, GPT-2): the reference's own Llama branch crashes."""
    import shutil

    from safetensors.torch import save_file

    cfg = json.load(open(os.path.join(source_dir, "config.json")))
    L = cfg.get("num_hidden_layers") or cfg.get("n_layer")
    if not L:
        raise ValueError(f"{source_dir}/config.json names no layer count (num_hidden_layers / n_layer)")
    tied = bool(cfg.get("tie_word_embeddings", cfg.get("model_type") == "gpt2"))
    ranges = shard_ranges(L, n_shards)
    base = os.path.join(output_dir, model_dir_name(model_name))
    tok_dir = os.path.join(base, "tokenizer")
    os.makedirs(tok_dir, exist_ok=True)
    for f in _TOKENIZER_FILES:
        if os.path.exists(os.path.join(source_dir, f)):
            shutil.copy2(os.path.join(source_dir, f), os.path.join(tok_dir, f))

    class _Out:  # one shard directory being written: a buffer of tensors flushed into numbered files
        def __init__(self, i):
            self.dir = os.path.join(base, f"shard_{i}")
            os.makedirs(self.dir, exist_ok=True)
            for stale in os.listdir(self.dir):
                if stale.endswith(".safetensors") or stale.endswith(".index.json"):
                    os.remove(os.path.join(self.dir, stale))
            self.buf, self.bytes, self.files, self.map, self.total = {}, 0, 0, {}, 0

        def add(self, name, t):
            nb = t.numel() * t.element_size()
            if self.buf and self.bytes + nb > max_file_bytes:
                self.flush()
            self.buf[name] = t.contiguous()
            self.bytes += nb
            self.total += nb

        def flush(self):
            if not self.buf:
                return
            self.files += 1
            fn = f"model-{self.files:05d}.safetensors"
            save_file(self.buf, os.path.join(self.dir, fn), metadata={"format": "pt"})
            for k in self.buf:
                self.map[k] = fn
            self.buf, self.bytes = {}, 0

        def close(self, i):
            self.flush()
            if self.files == 1:  # the usual single-file layout
                os.replace(os.path.join(self.dir, "model-00001.safetensors"), os.path.join(self.dir, "model.safetensors"))
            else:
                json.dump({"metadata": {"total_size": self.total}, "weight_map": self.map},
                          open(os.path.join(self.dir, "model.safetensors.index.json"), "w"), indent=2)
            json.dump(cfg, open(os.path.join(self.dir, "config.json"), "w"), indent=2)
            json.dump(shard_metadata(model_name, i, n_shards, L), open(os.path.join(self.dir, "metadata.json"), "w"), indent=2)
            gp = os.path.join(source_dir, "generation_config.json")
            if os.path.exists(gp):
                shutil.copy2(gp, os.path.join(self.dir, "generation_config.json"))

    outs = [_Out(i) for i in range(n_shards)]
    for name, load in _iter_source_tensors(source_dir):
        takers = [i for i, (a, b) in enumerate(ranges) if tensor_in_range(name, a, b, a == 0, b == L - 1, tied)]
        if not takers:
            continue  # rotary caches, attention-mask buffers ...: nothing a stage reads
        t = load()
        for i in takers:
            outs[i].add(name, t)
    for i, o in enumerate(outs):
        o.close(i)
    return [o.dir for o in outs]


def read_metadata(shard_path: str, model_name: str, shard_id) -> Dict[str, object]:
    p = os.path.join(shard_path, "metadata.json")
    if os.path.exists(p):
        return json.load(open(p))
    return default_metadata(model_name, shard_id)


def main(argv=None) -> int:
    """python -m dli_b200.sharding --model_name NAME --num_shards N --output_dir DIR
    (the arguments of the reference's `manage.py shard_model`, shard_model.py:11-14).  NAME is a local
    `save_pretrained` directory, or a name looked up under --cache_dir / $MODEL_CACHE_DIR (the hub is not reachable
    from the worker boxes, so nothing is downloaded)."""
    import argparse

    ap = argparse.ArgumentParser(description="Split a Hugging Face model directory into layer-range shard directories")
    ap.add_argument("--model_name", required=True, help="Name or path of the model to shard")
    ap.add_argument("--num_shards", type=int, default=2, help="Number of shards to create")
    ap.add_argument("--output_dir", default="model_shards", help="Directory to save the shards")
    ap.add_argument("--cache_dir", default=os.environ.get("MODEL_CACHE_DIR", "/app/model_cache"))
    ap.add_argument("--max_file_gb", type=float, default=2.0, help="largest output file (bounds the RAM used)")
    a = ap.parse_args(argv)
    src = None
    for cand in (a.model_name, os.path.join(a.cache_dir, model_dir_name(a.model_name)), os.path.join(a.cache_dir, a.model_name)):
        if os.path.isdir(cand) and os.path.exists(os.path.join(cand, "config.json")):
            src = cand
            break
    if src is None:
        print(f"Error: {a.model_name} is not a local model directory (looked in {a.cache_dir}); the Hugging Face hub is "
              "not reachable from here", flush=True)
        return 1
    name = a.model_name if not os.path.isdir(a.model_name) else os.path.basename(os.path.normpath(a.model_name))
    print(f"Sharding model {name} into {a.num_shards} shards...")
    paths = shard_model_dir(src, name, a.num_shards, a.output_dir, max_file_bytes=int(a.max_file_gb * (1 << 30)))
    for i, p in enumerate(paths):
        md = json.load(open(os.path.join(p, "metadata.json")))
        print(f"Shard {i}: layers {md['start_layer']} to {md['end_layer']} -> {p}")
    print(f"Successfully sharded model {name} into {a.num_shards} shards")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
