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
