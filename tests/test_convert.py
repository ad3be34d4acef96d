"""Real-weights importer: `python -m whisper_tflite_b200.convert` reads an HF
`transformers` (.safetensors / torch state dict) or openai-whisper (.pt) checkpoint, infers the architecture from the
tensor shapes and writes the engine's weight file; `--vocab-json` builds the vocab + filters blob the reference's Reader
parses (whisper.h:214-235). No real checkpoint is on this box: the files are written here from seeded weights in the
exact formats the two libraries use."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from whisper_tflite_b200 import convert, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _hf_safetensors(tmp, dims, W):
    from safetensors.numpy import save_file
    path = os.path.join(tmp, "model.safetensors")
    sd = {k: np.ascontiguousarray(v) for k, v in convert.canonical_to_hf(W).items()}
    del sd["proj_out.weight"]      # HF checkpoints do not store the tied projection
    save_file(sd, path)
    return path


def test_cli_safetensors_and_openai_pt_round_trip(tmp_path):
    import torch
    dims = synth.DIMS["micro"]
    W = synth.init_weights(dims, seed=3)
    st = _hf_safetensors(str(tmp_path), dims, W)
    out = str(tmp_path / "from_hf.b200w")
    r = subprocess.run([sys.executable, "-m", "whisper_tflite_b200.convert", st, out], cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "d_model 128, 2 heads, 2+2 layers" in r.stdout
    d2, W2 = synth.read_weight_file(out)
    assert (d2.d_model, d2.n_heads, d2.n_enc_layers, d2.n_dec_layers, d2.n_vocab, d2.n_text_ctx) == (128, 2, 2, 2, 51865, 448)
    for k in W:
        np.testing.assert_array_equal(W[k], W2[k])
    # openai-whisper layout: {"dims": ..., "model_state_dict": canonical names}
    pt = str(tmp_path / "micro.pt")
    torch.save({"dims": {"n_mels": 80}, "model_state_dict": {k: torch.from_numpy(v.copy()) for k, v in W.items()}}, pt)
    out2 = str(tmp_path / "from_openai.b200w")
    assert convert.main([pt, out2]) == 0
    assert open(out, "rb").read() == open(out2, "rb").read()
    # a checkpoint with a wrong shape is refused
    bad = dict(W)
    bad["encoder.ln_post.bias"] = np.zeros(7, np.float32)
    torch.save({"model_state_dict": {k: torch.from_numpy(np.array(v)) for k, v in bad.items()}}, pt)
    with pytest.raises(ValueError):
        convert.convert_checkpoint(pt, out2)


def test_vocab_json_to_blob_read_by_the_reference_reader(tmp_path):
    """vocab.json (GPT-2 byte-level spelling) -> blob; parsed back by the REFERENCE's own Reader (oracle/_ref) when it
    was built here, and by the host mirror's Reader."""
    b2u = {v: k for k, v in convert._bytes_to_unicode().items()}
    toks = [(" t%d" % i).encode() for i in range(50257)]
    toks[5], toks[6] = "é ü".encode(), bytes([0xF0, 0x9F])     # multi-byte UTF-8 and a split code point
    vj = str(tmp_path / "vocab.json")
    json.dump({"".join(b2u[b] for b in t): i for i, t in enumerate(toks)}, open(vj, "w", encoding="utf-8"))
    assert convert.vocab_json_to_tokens(vj) == toks
    assert convert.vocab_json_to_tokens(vj, multilingual=False) == toks[:50256]
    dims = synth.DIMS["micro"]
    st = _hf_safetensors(str(tmp_path), dims, synth.init_weights(dims, seed=1))
    blob = str(tmp_path / "fv.bin")
    assert convert.main([st, str(tmp_path / "w.b200w"), "--vocab-json", vj, "--vocab-out", blob]) == 0
    from whisper_tflite_b200 import build, host
    build.build_host()
    host.load()
    b = host.Blob(open(blob, "rb").read(), True)
    assert b.token(5) == toks[5] and b.token(6) == toks[6] and b.token(50256) == toks[50256]
    assert b.decode([5, 100, 50257]) == toks[5] + toks[100] + b.token(50257)
    np.testing.assert_array_equal(b.filters, synth.mel_filters())
    b.close()


@pytest.mark.gpu
def test_engine_from_converted_safetensors_gives_identical_ids(tmp_path):
    from tests.gpu_util import weight_file
    from whisper_tflite_b200 import capi
    path, dims, W = weight_file("tiny")
    st = _hf_safetensors(str(tmp_path), dims, W)
    out = str(tmp_path / "tiny.b200w")
    assert convert.main([st, out]) == 0
    pcm = synth.synth_pcm(3, seed=2)
    a = capi.Engine(path, max_batch=3)
    b = capi.Engine(out, max_batch=3)
    ia, la = a.transcribe_tokens(pcm)
    ib, lb = b.transcribe_tokens(pcm)
    np.testing.assert_array_equal(ia, ib)
    np.testing.assert_array_equal(la, lb)
    assert a.debug_hash("enc_out", 3) == b.debug_hash("enc_out", 3)
    a.close()
    b.close()
