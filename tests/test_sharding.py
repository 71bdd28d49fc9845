"""Shard directories: layout, metadata and selective loading (CPU).  The formula and the
metadata.json contents are checked against what the reference's own sharder wrote
(tests/golden/shard_ranges.json)."""
import json
import os

import pytest
import torch

from golden_util import GOLD, load_case, spec_kwargs
from dli_b200 import sharding
from oracle import decoder_ref as R


def _tiny():
    from dli_b200.engine import ModelSpec

    c = load_case("llama_tiny")
    cfg = spec_kwargs(c["config"])
    spec = ModelSpec.from_hf_config(cfg)
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"])
    return cfg, spec, sd


def test_metadata_equals_reference_sharder_output():
    for case in json.load(open(os.path.join(GOLD, "shard_ranges.json"))):
        L, N = case["n_layers"], case["n_shards"]
        assert sharding.shard_ranges(L, N) == [(m["start_layer"], m["end_layer"]) for m in case["metadata"]]
        for i, want in enumerate(case["metadata"]):
            assert sharding.shard_metadata(want["model_name"], i, N, L) == want
        assert sharding.model_dir_name("org/tiny") == case["dir"]


def test_write_shards_layout_and_selective_load(tmp_path):
    cfg, spec, sd = _tiny()
    paths = sharding.write_shards(sd, cfg, "org/tiny", 3, str(tmp_path))
    base = tmp_path / "org_tiny"
    assert [os.path.relpath(p, base) for p in paths] == ["shard_0", "shard_1", "shard_2"]
    assert (base / "tokenizer").is_dir()  # sibling tokenizer dir, as worker/app.py:168 expects
    seen = set()
    for i, p in enumerate(paths):
        md = json.load(open(os.path.join(p, "metadata.json")))
        assert md == sharding.shard_metadata("org/tiny", i, 3, spec.n_layers)
        part = sharding.load_hf_state_dict(p, md["start_layer"], md["end_layer"])
        layers = {int(k.split(".")[2]) for k in part if k.startswith("model.layers.")}
        assert layers == set(range(md["start_layer"], md["end_layer"] + 1))
        assert ("model.embed_tokens.weight" in part) == (i == 0)
        assert ("lm_head.weight" in part) == (i == 2) and ("model.norm.weight" in part) == (i == 2)
        for k, v in part.items():
            assert torch.equal(v, sd[k])
        seen |= set(part)
    assert seen == set(sd)  # every tensor stored exactly where it is needed, nothing lost
    # a shard directory holds only its share of the bytes (the reference stores a full model per shard)
    sizes = [os.path.getsize(os.path.join(p, "model.safetensors")) for p in paths]
    assert sum(sizes) < 1.05 * sum(v.numel() * v.element_size() for v in sd.values())


def test_reference_style_full_depth_directory_only_yields_in_range_tensors(tmp_path):
    """A reference shard directory is a full-depth save_pretrained() model whose out-of-range
    layers are random: the loader must read only [start, end]."""
    cfg, spec, sd = _tiny()
    from safetensors.torch import save_file

    poisoned = {k: (torch.full_like(v, 7.0) if k.startswith("model.layers.0.") else v) for k, v in sd.items()}
    save_file({k: v.contiguous() for k, v in poisoned.items()}, str(tmp_path / "model.safetensors"))
    json.dump(cfg, open(tmp_path / "config.json", "w"))
    part = sharding.load_hf_state_dict(str(tmp_path), 2, 3)
    assert not any(k.startswith(("model.layers.0.", "model.layers.1.", "model.embed_tokens")) for k in part)
    assert "lm_head.weight" in part and all(torch.equal(v, sd[k]) for k, v in part.items())
    # .bin format too
    os.remove(tmp_path / "model.safetensors")
    torch.save(sd, str(tmp_path / "pytorch_model.bin"))
    part = sharding.load_hf_state_dict(str(tmp_path), 0, 0)
    assert set(k for k in part if k.startswith("model.layers.")) == {k for k in sd if k.startswith("model.layers.0.")}
    assert "model.embed_tokens.weight" in part and "lm_head.weight" not in part


def test_gpt2_tied_head_reads_embedding_on_last_stage():
    assert sharding.tensor_in_range("transformer.wte.weight", 6, 11, False, True)
    assert not sharding.tensor_in_range("transformer.wte.weight", 3, 5, False, False)
    assert not sharding.tensor_in_range("transformer.wpe.weight", 6, 11, False, True)
    assert sharding.tensor_in_range("transformer.ln_f.bias", 6, 11, False, True)
    assert sharding.tensor_in_range("model.embed_tokens.weight", 2, 3, False, True, tied=True)
    assert not sharding.tensor_in_range("model.embed_tokens.weight", 2, 3, False, True, tied=False)
    with pytest.raises(FileNotFoundError):
        sharding._tensor_files("/nonexistent-dir")


@pytest.mark.parametrize("case_name,n_shards", [("llama_tiny", 3), ("gpt2_tiny", 2)])
def test_sharder_command_streams_a_model_directory(tmp_path, case_name, n_shards):
    """`python -m dli_b200.sharding --model_name --num_shards --output_dir` (the reference command's arguments,
    shard_model.py:11-14) on a multi-file source directory: same layout and metadata as the reference's sharder, the
    same tensors as write_shards(), written a few tensors at a time (tiny --max_file_gb forces multi-file shards)."""
    import subprocess
    import sys

    from safetensors.torch import save_file

    from dli_b200.engine import ModelSpec

    c = load_case(case_name)
    cfg = spec_kwargs(c["config"])
    spec = ModelSpec.from_hf_config(cfg)
    sd = R.seeded_state_dict(R.ModelSpec(**spec.__dict__), c["seed"], c["std"])
    src = tmp_path / "cache" / "org_tiny"
    src.mkdir(parents=True)
    names = sorted(sd)
    half = len(names) // 2
    files = {"model-00001-of-00002.safetensors": names[:half], "model-00002-of-00002.safetensors": names[half:]}
    for fn, ks in files.items():
        save_file({k: sd[k].contiguous() for k in ks}, str(src / fn))
    json.dump({"metadata": {}, "weight_map": {k: fn for fn, ks in files.items() for k in ks}},
              open(src / "model.safetensors.index.json", "w"))
    json.dump(cfg, open(src / "config.json", "w"))
    json.dump({"eos_token_id": 2}, open(src / "generation_config.json", "w"))
    (src / "tokenizer.json").write_text("{}")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = tmp_path / "model_shards"
    p = subprocess.run([sys.executable, "-m", "dli_b200.sharding", "--model_name", "org/tiny", "--num_shards", str(n_shards),
                        "--output_dir", str(out), "--cache_dir", str(tmp_path / "cache"), "--max_file_gb", "0.00002"],
                       cwd=root, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout + p.stderr
    assert f"Successfully sharded model org/tiny into {n_shards} shards" in p.stdout
    base = out / "org_tiny"
    assert (base / "tokenizer" / "tokenizer.json").exists()
    want_paths = sharding.write_shards(sd, cfg, "org/tiny", n_shards, str(tmp_path / "ref"))
    seen = set()
    for i in range(n_shards):
        d = base / f"shard_{i}"
        md = json.load(open(d / "metadata.json"))
        assert md == sharding.shard_metadata("org/tiny", i, n_shards, spec.n_layers)
        assert json.load(open(d / "config.json")) == cfg and (d / "generation_config.json").exists()
        assert (d / "model.safetensors.index.json").exists()  # several small files + an index
        got = sharding.load_hf_state_dict(str(d), md["start_layer"], md["end_layer"])
        want = sharding.load_hf_state_dict(want_paths[i], md["start_layer"], md["end_layer"])
        assert set(got) == set(want) and all(torch.equal(got[k], want[k]) for k in want)
        seen |= set(got)
    assert seen == set(sd)
    # a name that is nowhere local is an error, not a download
    p = subprocess.run([sys.executable, "-m", "dli_b200.sharding", "--model_name", "org/absent", "--cache_dir", str(tmp_path)],
                       cwd=root, capture_output=True, text=True, timeout=120)
    assert p.returncode == 1 and "not a local model directory" in p.stdout
