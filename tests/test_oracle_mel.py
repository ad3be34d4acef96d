"""Pins the CPU mel oracle: restatement (oracle/mel_oracle.c) == the reference's own compiled code
(oracle/_ref) bit for bit, and both == tests/golden/mel_golden.npz (generated from the reference)."""
import os

import numpy as np
import pytest

from tests.golden.make_mel_golden import STRIDE, golden_inputs
from whisper_tflite_b200 import synth

GOLD = os.path.join(os.path.dirname(__file__), "golden", "mel_golden.npz")


def test_restatement_matches_golden(oracle_built, filters):
    g = np.load(GOLD)
    pcm = golden_inputs()
    for c in range(pcm.shape[0]):
        m = oracle_built.oracle_log_mel(pcm[c], filters, n_threads=8)
        assert m.shape == (80, 3000)
        np.testing.assert_array_equal(m[:, ::STRIDE], g["sample"][c])
        assert m.max() == g["mx"][c] and m.min() == g["mn"][c]
        assert abs(np.abs(m).sum(dtype=np.float64) - g["abs_sum"][c]) <= 1e-9 * max(1.0, g["abs_sum"][c])


def test_silence_and_clamp(oracle_built, filters):
    # all-zero PCM: every bin hits the 1e-10 clamp -> log10 = -10 -> (-10+4)/4 (whisper.cpp:176-213)
    m = oracle_built.oracle_log_mel(np.zeros(synth.N_SAMPLES, np.float32), filters, 4)
    assert np.all(m == np.float32(-1.5))


def test_restatement_equals_compiled_reference(oracle_built, filters):
    if not oracle_built.have_ref():
        pytest.skip("oracle/_ref not built (reference tree absent)")
    rs = np.random.RandomState(11)
    for n in (400, 200, 100, 50, 25, 7):
        x = rs.randn(n).astype(np.float32)
        np.testing.assert_array_equal(oracle_built.ref_fft(x), oracle_built.oracle_fft(x))
    pcm = synth.synth_pcm(1, seed=3)[0]
    np.testing.assert_array_equal(oracle_built.ref_log_mel(pcm, filters, 0),
                                  oracle_built.oracle_log_mel(pcm, filters, 8))
    # ragged: shorter input -> n_len = n/160 frames, zero-filled past the end (whisper.cpp:123,149-153)
    short = pcm[:160 * 37 + 55]
    a = oracle_built.ref_log_mel(short, filters, 3)
    b = oracle_built.oracle_log_mel(short, filters, 2)
    assert a.shape == (80, 37)
    np.testing.assert_array_equal(a, b)


def test_oracle_close_to_float64_truth(oracle_built, filters):
    """The reference's f32 recursive FFT sits ~1e-4 (abs, range ~2) from an f64 evaluation of the same
    definition; this bounds how close ANY correct implementation can be asked to match it."""
    pcm = synth.synth_pcm(1, seed=0)[0].astype(np.float64)
    hann = (0.5 * (1.0 - np.cos(2.0 * np.pi * np.arange(400) / 400))).astype(np.float32).astype(np.float64)
    x = np.concatenate([pcm, np.zeros(400)])
    frames = np.lib.stride_tricks.sliding_window_view(x, 400)[::160][:3000] * hann
    p = np.abs(np.fft.rfft(frames, axis=1)) ** 2
    p[:, 1:200] *= 2.0  # mirrored-bin fold, whisper.cpp:164-166
    m = np.log10(np.maximum(p @ filters.astype(np.float64).T, 1e-10)).T
    m = (np.maximum(m, m.max() - 8.0) + 4.0) / 4.0
    got = oracle_built.oracle_log_mel(pcm.astype(np.float32), filters, 8)
    assert np.abs(got - m).max() < 2e-4 * (m.max() - m.min())


def test_argmax_last_index(oracle_built):
    x = np.array([1, 5, 3, 5, 2, 5, 0], np.float32)
    assert oracle_built.oracle_argmax_last(x) == 5
    assert oracle_built.oracle_argmax_last(np.array([2, 1], np.float32)) == 0
