import hashlib
import os

import numpy as np
import pytest

from whisper_tflite_b200 import convert, synth


def test_mel_filters_pinned():
    f = synth.mel_filters()
    assert f.shape == (80, 201) and f.dtype == np.float32
    # sha256 of HF WhisperFeatureExtractor().mel_filters.T.astype(f32) (== OpenAI mel_filters.npz mel_80)
    assert hashlib.sha256(f.tobytes()).hexdigest() == \
        "2150c30cbbeb6029f52002ffa666c1c72d83dbf53f463cb8462052055806e891"
    assert int((f != 0).sum()) == 391


def test_mel_filters_equal_hf():
    transformers = pytest.importorskip("transformers")
    m = transformers.WhisperFeatureExtractor().mel_filters
    np.testing.assert_array_equal(synth.mel_filters(), m.T.astype(np.float32))


def test_pcm_deterministic_and_bounded():
    a = synth.synth_pcm(3, seed=0)
    b = synth.synth_pcm(5, seed=0)
    np.testing.assert_array_equal(a, b[:3])
    assert a.shape == (3, 480000) and np.abs(a).max() <= 1.0
    assert (a[:, -8:] == 0).any()


def test_weight_file_roundtrip(tmp_path):
    dims = synth.DIMS["micro"]
    W = synth.init_weights(dims, seed=0)
    p = str(tmp_path / "m.b200w")
    synth.write_weight_file(p, dims, W)
    d2, W2 = synth.read_weight_file(p)
    assert (d2.d_model, d2.n_heads, d2.n_enc_layers, d2.n_dec_layers, d2.n_vocab) == (128, 2, 2, 2, 51865)
    assert set(W2) == set(W)
    for k in W:
        np.testing.assert_array_equal(W[k], W2[k])


def test_hf_name_mapping_roundtrip():
    dims = synth.DIMS["micro"]
    W = synth.init_weights(dims, seed=0)
    hf = convert.canonical_to_hf(W)
    back = convert.hf_to_canonical(hf, dims)
    assert set(back) == set(W)
    for k in W:
        np.testing.assert_array_equal(W[k], back[k])


def test_vocab_blob_against_reference_reader(tmp_path, oracle_built):
    if not oracle_built.have_ref():
        pytest.skip("oracle/_ref not built")
    toks = synth.synth_vocab_tokens(50257)
    p = str(tmp_path / "filters_vocab_multilingual.bin")
    size = synth.write_vocab_blob(p, synth.mel_filters(), toks)
    assert os.path.getsize(p) == size
    blob = open(p, "rb").read()
    rv = oracle_built.RefVocab(blob, multilingual=True)
    assert rv.n_vocab == 51865
    np.testing.assert_array_equal(rv.filters, synth.mel_filters())
    assert rv.special["eot"] == 50257 and rv.special["sot"] == 50258 and rv.special["not"] == 50363
    assert rv.special["transcribe"] == 50359  # NOT shifted (whisper.cpp:218-226)
    assert rv.token(17) == b" t17"
    assert rv.token(50261) == b"<|lang-de|>"
