"""Pins the numpy encoder/decoder oracle against tests/golden/model_golden_*.npz, which were produced
by HF transformers (an independent implementation of the same architecture) on the same seeded
weights — see tests/golden/make_model_golden.py. The TFLite boundary itself stays UNPINNED
(no TFLite, no .tflite, no reference golden on this box; oracle/model_ref.py header)."""
import os

import numpy as np
import pytest

from oracle import model_ref
from whisper_tflite_b200 import synth

HERE = os.path.dirname(__file__)


@pytest.mark.parametrize("size", ["micro", "tiny"])
def test_oracle_matches_hf_golden(size, oracle_built, filters):
    g = np.load(os.path.join(HERE, "golden", f"model_golden_{size}.npz"))
    dims = synth.DIMS[size]
    W = synth.init_weights(dims, seed=int(g["seed"]))
    pcm = synth.synth_pcm(int(g["n_clips"]), seed=int(g["seed"]))
    n_check = 1 if size == "tiny" else int(g["n_clips"])
    for c in range(n_check):
        mel = oracle_built.oracle_log_mel(pcm[c], filters, 8)
        xa = model_ref.encoder_forward(W, dims, mel)
        ref = g["enc_sample"][c]
        got = xa[::25, ::(dims.d_model // 32)]
        assert np.abs(got - ref).max() <= 2e-5 * np.abs(ref).max()
        ids, rows = model_ref.greedy_decode(W, dims, xa, list(g["prompt"]), int(g["eot"]), return_logits=True)
        n = int(g["lens"][c])
        assert ids == list(g["ids"][c][:n])
        for s in range(n - 4):
            np.testing.assert_allclose(rows[s][g["top_idx"][c][s]], g["top_val"][c][s], rtol=0, atol=5e-5)


def test_prompt_and_eot_stop():
    assert model_ref.reference_prompt(True) == [50258, 50261, 50359, 50363]  # whisper.cpp:331
    dims = synth.DIMS["micro"]
    W = synth.init_weights(dims, seed=1)
    xa = np.random.RandomState(0).randn(1500, dims.d_model).astype(np.float32)
    full = model_ref.greedy_decode(W, dims, xa, model_ref.reference_prompt(True), eot=50257)
    assert len(full) == 31  # 4 prompt + 27 new (whisper.cpp:364-367)
    # declare the 3rd generated token to be EOT: the loop must stop right after appending it
    eot = full[6]
    cut = model_ref.greedy_decode(W, dims, xa, model_ref.reference_prompt(True), eot=eot)
    assert cut == full[:full.index(eot, 4) + 1]


def test_argmax_tie_takes_last():
    assert model_ref.argmax_last(np.array([0.0, 2.0, 2.0, 1.0], np.float32)) == 2
