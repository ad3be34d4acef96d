"""Weight-name mapping between HF `transformers` Whisper checkpoints and the engine's canonical
(openai-whisper) names — the first half of SURVEY.md §8(f) rank 1 (real-weights importer).

`hf_to_canonical(state_dict)` accepts any mapping name -> array-like (torch tensors or numpy) and
returns the dict `synth.write_weight_file` expects; `canonical_to_hf` is its inverse (used by
tests/golden/make_model_golden.py to load seeded weights into the HF model).
"""
from __future__ import annotations

import re
from typing import Dict

import numpy as np

_ATTN = {"q_proj": "query", "k_proj": "key", "v_proj": "value", "out_proj": "out"}


def _hf_name(canon: str) -> str:
    m = re.match(r"(encoder|decoder)\.blocks\.(\d+)\.(.*)", canon)
    if m:
        side, i, rest = m.groups()
        pre = f"model.{side}.layers.{i}."
        for hf, ca in _ATTN.items():
            if rest.startswith("attn." + ca + "."):
                return pre + "self_attn." + hf + rest[len("attn." + ca):]
            if rest.startswith("cross_attn." + ca + "."):
                return pre + "encoder_attn." + hf + rest[len("cross_attn." + ca):]
        table = {"attn_ln": "self_attn_layer_norm", "cross_attn_ln": "encoder_attn_layer_norm",
                 "mlp_ln": "final_layer_norm", "mlp.0": "fc1", "mlp.2": "fc2"}
        for ca, hf in table.items():
            if rest.startswith(ca + "."):
                return pre + hf + rest[len(ca):]
        raise KeyError(canon)
    table = {
        "encoder.conv1.": "model.encoder.conv1.", "encoder.conv2.": "model.encoder.conv2.",
        "encoder.positional_embedding": "model.encoder.embed_positions.weight",
        "encoder.ln_post.": "model.encoder.layer_norm.",
        "decoder.token_embedding.weight": "model.decoder.embed_tokens.weight",
        "decoder.positional_embedding": "model.decoder.embed_positions.weight",
        "decoder.ln.": "model.decoder.layer_norm.",
    }
    for ca, hf in table.items():
        if canon.startswith(ca):
            return hf + canon[len(ca):]
    raise KeyError(canon)


def canonical_to_hf(weights: Dict[str, np.ndarray]) -> Dict[str, np.ndarray]:
    out = {_hf_name(k): v for k, v in weights.items()}
    out["proj_out.weight"] = weights["decoder.token_embedding.weight"]  # tied (modeling_whisper.py:966-971)
    return out


def hf_to_canonical(state_dict, dims) -> Dict[str, np.ndarray]:
    from .synth import weight_names

    def to_np(x):
        if hasattr(x, "detach"):
            x = x.detach().float().cpu().numpy()
        return np.ascontiguousarray(x, dtype=np.float32)

    out = {}
    for name, shape in weight_names(dims):
        a = to_np(state_dict[_hf_name(name)])
        if tuple(a.shape) != tuple(shape):
            raise ValueError(f"{name}: expected {shape}, checkpoint has {a.shape}")
        out[name] = a
    return out
