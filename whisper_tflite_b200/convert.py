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


# ----------------------------------------------------------------------------------------------
# Checkpoint files -> engine weight file, and the vocab + filters blob
#
#   python -m whisper_tflite_b200.convert model.safetensors whisper-tiny.b200w
#   python -m whisper_tflite_b200.convert tiny.pt whisper-tiny.b200w --vocab-json vocab.json --vocab-out filters_vocab_multilingual.bin
#
# Accepted inputs: an HF `transformers` Whisper checkpoint (.safetensors, or a torch .bin/.pt state dict with
# `model.encoder...` names) or an openai-whisper checkpoint (.pt holding {"dims", "model_state_dict"} with the canonical
# names the reference's export uses: export/generate_onnx.py:85-130). The architecture is read off the tensor shapes.
# ----------------------------------------------------------------------------------------------
def load_checkpoint(path: str):
    """-> (state dict name -> numpy/torch array, 'hf' | 'openai')."""
    if path.endswith(".safetensors"):
        from safetensors.numpy import load_file
        sd = load_file(path)
    else:
        import torch
        obj = torch.load(path, map_location="cpu", weights_only=True)
        sd = obj.get("model_state_dict", obj) if isinstance(obj, dict) else obj
    kind = "hf" if any(k.startswith("model.encoder.") for k in sd) else "openai"
    return sd, kind


def infer_dims(sd, kind: str):
    """WhisperDims from tensor shapes (head_dim is 64 in every Whisper size: SURVEY.md appendix A.4)."""
    from .synth import WhisperDims
    name = (lambda c: _hf_name(c)) if kind == "hf" else (lambda c: c)
    shp = lambda c: tuple(sd[name(c)].shape)
    d, n_mels, _ = shp("encoder.conv1.weight")
    n_audio_ctx = shp("encoder.positional_embedding")[0]
    n_vocab = shp("decoder.token_embedding.weight")[0]
    n_text_ctx = shp("decoder.positional_embedding")[0]

    def count(side):
        i = 0
        while name(f"{side}.blocks.{i}.attn_ln.weight") in sd:
            i += 1
        return i

    return WhisperDims("converted", d_model=d, n_heads=d // 64, n_enc_layers=count("encoder"), n_dec_layers=count("decoder"),
                       n_vocab=n_vocab, n_mels=n_mels, n_audio_ctx=n_audio_ctx, n_text_ctx=n_text_ctx)


def convert_checkpoint(src: str, dst: str):
    """Checkpoint file -> <prefix>.b200w. Returns the inferred dims."""
    from .synth import weight_names, write_weight_file
    sd, kind = load_checkpoint(src)
    dims = infer_dims(sd, kind)
    if kind == "hf":
        W = hf_to_canonical(sd, dims)
    else:
        W = {}
        for n, shape in weight_names(dims):
            a = sd[n]
            a = a.detach().float().cpu().numpy() if hasattr(a, "detach") else np.asarray(a)
            if tuple(a.shape) != tuple(shape):
                raise ValueError(f"{n}: expected {shape}, checkpoint has {a.shape}")
            W[n] = np.ascontiguousarray(a, dtype=np.float32)
    write_weight_file(dst, dims, W)
    return dims


def _bytes_to_unicode():
    """GPT-2's printable stand-ins for the 256 byte values (the table HF/openai vocab.json files are written in)."""
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAC + 1)) + list(range(0xAE, 0xFF + 1))
    cs = bs[:]
    n = 0
    for b in range(256):
        if b not in bs:
            bs.append(b)
            cs.append(256 + n)
            n += 1
    return {chr(c): b for b, c in zip(bs, cs)}


def vocab_json_to_tokens(vocab_json: str, multilingual: bool = True):
    """HF `vocab.json` (token string -> id, byte-level BPE spelling) -> the raw byte string of every id below the
    special tokens: ids [0, 50257) for the multilingual vocabulary, [0, 50256) for the English one — what the
    reference's blob stores before Reader synthesises the special-token names (whisper.cpp:577-604)."""
    import json
    table = json.load(open(vocab_json, encoding="utf-8"))
    inv = _bytes_to_unicode()
    n = 50257 if multilingual else 50256
    toks = [b""] * n
    for s, i in table.items():
        if i < n:
            toks[i] = bytes(inv[ch] for ch in s)
    return toks


def main(argv=None):
    import argparse
    from . import synth
    ap = argparse.ArgumentParser(prog="python -m whisper_tflite_b200.convert", description=__doc__)
    ap.add_argument("checkpoint", help=".safetensors / .bin / .pt (HF transformers or openai-whisper)")
    ap.add_argument("out", help="engine weight file to write (<prefix>.b200w, next to which EncDec/Monolith look for it)")
    ap.add_argument("--vocab-json", help="HF vocab.json: also write the vocab + mel-filter blob (whisper.h:214-235)")
    ap.add_argument("--vocab-out", default="filters_vocab_multilingual.bin")
    ap.add_argument("--english", action="store_true", help="English-only vocabulary (50256 ordinary tokens)")
    a = ap.parse_args(argv)
    dims = convert_checkpoint(a.checkpoint, a.out)
    print(f"wrote {a.out}: d_model {dims.d_model}, {dims.n_heads} heads, {dims.n_enc_layers}+{dims.n_dec_layers} layers, "
          f"vocab {dims.n_vocab}, text ctx {dims.n_text_ctx}")
    if a.vocab_json:
        size = synth.write_vocab_blob(a.vocab_out, synth.mel_filters(), vocab_json_to_tokens(a.vocab_json, not a.english))
        print(f"wrote {a.vocab_out}: {size} bytes")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
