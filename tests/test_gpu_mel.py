"""Parity of the CUDA log-mel frontend (through the C ABI) with the reference's log_mel_spectrogram.
Tolerance (north_star): 1e-4 relative, read as max-abs <= 1e-4 x output range (range ~2.0) — the
reference's own f32 recursive FFT sits 7.5e-5 from an f64 evaluation."""
import os

import numpy as np
import pytest

from tests.golden.make_mel_golden import STRIDE, golden_inputs
from whisper_tflite_b200 import synth

pytestmark = pytest.mark.gpu
TOL = 1e-4


@pytest.fixture(scope="module")
def eng():
    from whisper_tflite_b200 import capi
    e = capi.Engine(None, max_batch=64)
    yield e
    e.close()


def test_mel_vs_golden_fixture(eng):
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "mel_golden.npz"))
    pcm = golden_inputs()
    out = eng.log_mel(pcm)
    assert out.shape == (6, 80, 3000)
    for c in range(6):
        rng = max(1.0, float(g["mx"][c] - g["mn"][c]))
        assert np.abs(out[c][:, ::STRIDE] - g["sample"][c]).max() <= TOL * rng
        assert abs(float(out[c].max()) - float(g["mx"][c])) <= TOL * rng
        assert abs(float(out[c].min()) - float(g["mn"][c])) <= TOL * rng
    assert np.all(out[2] == np.float32(-1.5))  # silence: every bin at the 1e-10 clamp


def test_mel_vs_oracle_full_arrays(eng, oracle_built, filters):
    pcm = synth.synth_pcm(4, seed=21)
    out = eng.log_mel(pcm)
    for c in range(4):
        ref = oracle_built.oracle_log_mel(pcm[c], filters, 8)
        assert np.abs(out[c] - ref).max() <= TOL * (ref.max() - ref.min())


def test_mel_vs_compiled_reference(eng, oracle_built, filters):
    if not oracle_built.have_ref():
        pytest.skip("oracle/_ref not present")
    pcm = synth.synth_pcm(1, seed=22)
    out = eng.log_mel(pcm)
    ref = oracle_built.ref_log_mel(pcm[0], filters, 0)
    assert np.abs(out[0] - ref).max() <= TOL * (ref.max() - ref.min())


@pytest.mark.parametrize("n_samples", [160, 160 * 31 + 159, 160 * 32, 160 * 70 + 13, 479999])
def test_mel_ragged_lengths(eng, oracle_built, filters, n_samples):
    pcm = synth.synth_pcm(2, seed=23)[:, :n_samples].copy()
    out = eng.log_mel(pcm)
    assert out.shape == (2, 80, n_samples // 160)
    for c in range(2):
        ref = oracle_built.oracle_log_mel(pcm[c], filters, 8)
        assert np.abs(out[c] - ref).max() <= TOL * max(1e-3, ref.max() - ref.min())


def test_mel_full_batch_properties(eng, oracle_built, filters):
    """BASELINE config-2 size (64 x 30 s): spot-check clips against the oracle, and size-independent
    properties on all of them: batch position independence, clamp floor = max - 8 -> (max-8+4)/4."""
    pcm = synth.synth_pcm(64, seed=0)
    out = eng.log_mel(pcm)
    for c in (0, 17, 63):
        ref = oracle_built.oracle_log_mel(pcm[c], filters, 8)
        assert np.abs(out[c] - ref).max() <= TOL * (ref.max() - ref.min())
    perm = np.random.RandomState(0).permutation(64)
    out2 = eng.log_mel(np.ascontiguousarray(pcm[perm]))
    np.testing.assert_array_equal(out2, out[perm])
    mx, mn = out.max(axis=(1, 2)), out.min(axis=(1, 2))
    assert np.all(mx - mn <= 2.0 + 1e-6)  # (max - (max-8)) / 4
    zero_tail = np.array([np.all(p[-80:] == 0) for p in pcm])
    assert np.allclose((mx - mn)[zero_tail], 2.0, atol=1e-6)


def test_mel_device_buffers(eng):
    torch = pytest.importorskip("torch")
    pcm = synth.synth_pcm(3, seed=5)
    host = eng.log_mel(pcm)
    d_pcm = torch.from_numpy(pcm).cuda()
    d_out = torch.empty(3, 80, 3000, device="cuda")
    eng.log_mel(d_pcm, out=d_out)
    eng.synchronize()
    np.testing.assert_array_equal(d_out.cpu().numpy(), host)


def test_mel_errors(eng):
    from whisper_tflite_b200 import capi
    with pytest.raises(capi.WhisperB200Error):
        eng.log_mel(np.zeros((65, 480000), np.float32))  # > max_batch
    with pytest.raises(capi.WhisperB200Error):
        eng.log_mel(np.zeros((1, 100), np.float32))  # < one hop


def test_int16_pcm_is_bit_equal_to_f32_path(filters):
    """wt_log_mel_i16: the staging load converts static_cast<float>(s) / static_cast<float>(INT16_MAX) (wav_util.cpp:70-73);
    the mel must be BIT-equal to wt_log_mel on the host-converted samples — full windows, a ragged length that is not
    a multiple of 8 (scalar staging path), host and device buffers."""
    import torch
    from whisper_tflite_b200 import capi
    rs = np.random.RandomState(8)
    eng = capi.Engine(None, max_batch=3, filters=filters)
    for n in (480000, 160 * 37 + 3):
        p16 = rs.randint(-32768, 32768, size=(3, n)).astype(np.int16)
        p16[1, n // 2:] = 0
        f32 = (p16.astype(np.float32) / np.float32(32767.0)).astype(np.float32)
        a = eng.log_mel(p16)
        b = eng.log_mel(f32)
        assert a.shape == b.shape == (3, 80, n // 160)
        np.testing.assert_array_equal(a.view(np.uint32), b.view(np.uint32))
        c = eng.log_mel(torch.from_numpy(p16).cuda())
        np.testing.assert_array_equal(c.view(np.uint32), b.view(np.uint32))
    eng.close()
