"""Load the committed fixtures of tests/golden/ (made by oracle/gen_golden.py)."""
import json
import os

import numpy as np
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MODEL_CASES = ["llama_tiny", "llama_hd128", "mistral_tiny", "gpt2_tiny"]


def _bf16(a):
    return torch.from_numpy(a.copy()).view(torch.bfloat16)


def load_case(name):
    z = np.load(os.path.join(GOLD, f"{name}.npz"))
    return {
        "config": json.loads(str(z["config"])),
        "seed": int(z["seed"]),
        "std": float(z["std"]),
        "prompts": torch.from_numpy(z["prompts"]),
        "tokens": torch.from_numpy(z["hf_bf16_tokens"]),               # [B, N]
        "logits_bf16": _bf16(z["hf_bf16_logits"]).float(),            # [N, B, V]
        "logits_fp32": torch.from_numpy(z["hf_fp32_logits"]).float(),  # [N, B, V] (stored as fp16)
        "margin": torch.from_numpy(z["hf_fp32_margin"]),               # [N, B]
        "hf_err": torch.from_numpy(z["hf_bf16_err"]),                  # [N, B]
        "boundary_hidden": _bf16(z["boundary_hidden"]).float(),        # [B, T, d]
        "boundary_layer": int(z["boundary_layer"]),
    }


def spec_kwargs(config):
    """config kwargs (as stored) -> dict with model_type for ModelSpec.from_hf_config."""
    c = dict(config)
    if c["model_type"] != "gpt2":
        c.setdefault("tie_word_embeddings", False)
    return c
