"""Logic checks of the SIMT CUDA kernels (csrc/*.cuh) compiled with g++ against tests/emu's CPU SIMT
emulator, compared with the oracle — so indexing/algorithm errors are caught without a GPU.
The real parity tests (through the C-ABI, on the B200) are tests/test_gpu_*.py."""
import ctypes
import os

import numpy as np
import pytest

from oracle import model_ref
from tests.emu.build import build
from tests.golden.make_mel_golden import STRIDE, golden_inputs
from whisper_tflite_b200 import synth

fp = ctypes.POINTER(ctypes.c_float)
ip = ctypes.POINTER(ctypes.c_int)
LL = ctypes.c_longlong


def P(a):
    return a.ctypes.data_as(fp)


@pytest.fixture(scope="module")
def mel_lib():
    return ctypes.CDLL(build("mel"))


@pytest.fixture(scope="module")
def simt():
    return ctypes.CDLL(build("simt"))


def emu_mel(lib, pcm, F):
    pcm = np.ascontiguousarray(pcm, np.float32)
    B, n = pcm.shape
    raw = np.zeros((B, F.shape[0], n // 160), np.float32)
    out = np.zeros_like(raw)
    lib.emu_log_mel(P(pcm), B, n, P(F), F.shape[0], P(raw), P(out))
    return raw, out


def test_mel_kernel_vs_oracle_ragged(mel_lib, oracle_built, filters):
    pcm = synth.synth_pcm(2, seed=4)[:, :160 * 70 + 13]  # 70 frames: 2 full tiles + ragged third
    raw, out = emu_mel(mel_lib, pcm, filters)
    for c in range(2):
        ref, ref_raw = oracle_built.oracle_log_mel(pcm[c], filters, 4, return_raw=True)
        assert out[c].shape == ref.shape == (80, 70)
        assert np.abs(out[c] - ref).max() <= 1e-4 * (ref.max() - ref.min())


def test_mel_kernel_vs_golden_edge_clips(mel_lib, filters):
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "mel_golden.npz"))
    pcm = golden_inputs()[2:5]  # silence, DC, impulse train (full 30 s)
    _, out = emu_mel(mel_lib, pcm, filters)
    for i, c in enumerate(range(2, 5)):
        rng = max(1e-3, float(g["mx"][c] - g["mn"][c]))
        assert np.abs(out[i][:, ::STRIDE] - g["sample"][c]).max() <= 1e-4 * max(rng, 1.0)


def test_mel_kernel_dense_filterbank(mel_lib, oracle_built):
    # a dense (non-sparse) bank exercises the generic packed-row path
    F = np.abs(np.random.RandomState(2).randn(80, 201)).astype(np.float32) * 0.01
    pcm = synth.synth_pcm(1, seed=9)[:, :160 * 32]
    _, out = emu_mel(mel_lib, pcm, F)
    ref = oracle_built.oracle_log_mel(pcm[0], F, 2)
    assert np.abs(out[0] - ref).max() <= 1e-4 * (ref.max() - ref.min())


def sgemm(simt, A, amap, W, bias, residual, pos, C, cmap, M, N, K, flags):
    simt.emu_sgemm(P(A), LL(amap[0]), amap[1], amap[2], P(W), P(bias) if bias is not None else None,
                   P(residual) if residual is not None else None, P(pos) if pos is not None else None,
                   pos.shape[0] if pos is not None else 1, P(C), LL(cmap[0]), cmap[1], cmap[2], M, N, K, flags)


def test_sgemm_plain_and_epilogues(simt):
    rs = np.random.RandomState(0)
    M, N, K = 150, 100, 70
    A = rs.randn(M, K).astype(np.float32)
    W = rs.randn(N, K).astype(np.float32)
    b = rs.randn(N).astype(np.float32)
    R = rs.randn(M, N).astype(np.float32)
    C = np.zeros((M, N), np.float32)
    sgemm(simt, A, (0, M, K), W, b, R, None, C, (0, M, N), M, N, K, 1 | 2 | 4)
    ref = model_ref.gelu((A @ W.T + b).astype(np.float32)) + R
    np.testing.assert_allclose(C, ref, rtol=1e-4, atol=1e-4)


def test_conv_as_overlapping_row_gemm(simt):
    """conv1 (k3,s1,p1)+GELU and conv2 (k3,s2,p1)+GELU+pos as GEMMs over a zero-padded time-major buffer."""
    rs = np.random.RandomState(1)
    B, T, Cin, Cout = 2, 40, 16, 24
    x = rs.randn(B, Cin, T).astype(np.float32)
    w1 = rs.randn(Cout, Cin, 3).astype(np.float32) * 0.2
    b1 = rs.randn(Cout).astype(np.float32)
    xt = np.zeros((B, T + 2, Cin), np.float32)
    xt[:, 1:-1] = x.transpose(0, 2, 1)
    w1p = np.ascontiguousarray(w1.transpose(0, 2, 1).reshape(Cout, 3 * Cin))  # [n][dk*Cin + c]
    y1 = np.zeros((B, T + 2, Cout), np.float32)
    # A row (b,t) = xt[b, t:t+3, :] flattened; C row (b,t) -> y1[b, t+1]
    sgemm(simt, xt, ((T + 2) * Cin, T, Cin), w1p, b1, None, None, y1.reshape(-1)[Cout:], ((T + 2) * Cout, T, Cout),
          B * T, Cout, 3 * Cin, 1 | 2)
    for b in range(B):
        ref = model_ref.gelu(model_ref.conv1d_k3(x[b], w1, b1, 1)).T
        np.testing.assert_allclose(y1[b, 1:-1], ref, rtol=1e-4, atol=1e-4)
    assert np.all(y1[:, 0] == 0) and np.all(y1[:, -1] == 0)
    w2 = rs.randn(Cout, Cout, 3).astype(np.float32) * 0.2
    b2 = rs.randn(Cout).astype(np.float32)
    pos = rs.randn(T // 2, Cout).astype(np.float32)
    w2p = np.ascontiguousarray(w2.transpose(0, 2, 1).reshape(Cout, 3 * Cout))
    y2 = np.zeros((B * (T // 2), Cout), np.float32)
    sgemm(simt, y1, ((T + 2) * Cout, T // 2, 2 * Cout), w2p, b2, None, pos, y2, (0, B * (T // 2), Cout),
          B * (T // 2), Cout, 3 * Cout, 1 | 2 | 8)
    for b in range(B):
        ref = model_ref.gelu(model_ref.conv1d_k3(y1[b, 1:-1].T.copy(), w2, b2, 2)).T + pos
        np.testing.assert_allclose(y2.reshape(B, T // 2, Cout)[b], ref, rtol=1e-4, atol=1e-4)


def test_layernorm(simt):
    rs = np.random.RandomState(2)
    x = rs.randn(37, 384).astype(np.float32) * 3 + 1
    g = rs.randn(384).astype(np.float32)
    b = rs.randn(384).astype(np.float32)
    y = np.zeros_like(x)
    simt.emu_layernorm(P(x), P(g), P(b), P(y), 37, 384)
    np.testing.assert_allclose(y, model_ref.layer_norm(x, g, b), rtol=1e-4, atol=1e-5)


def test_encoder_attention(simt):
    rs = np.random.RandomState(3)
    B, S, H = 2, 200, 2
    D = H * 64
    qkv = rs.randn(B, S, 3 * D).astype(np.float32)
    out = np.zeros((B, S, D), np.float32)
    simt.emu_attention(P(qkv), P(out), B, S, H)
    for b in range(B):
        q, k, v = qkv[b, :, :D], qkv[b, :, D:2 * D], qkv[b, :, 2 * D:]
        qh = q.reshape(S, H, 64).transpose(1, 0, 2) * 0.125
        kh = k.reshape(S, H, 64).transpose(1, 2, 0)
        vh = v.reshape(S, H, 64).transpose(1, 0, 2)
        ref = (model_ref.softmax(qh @ kh) @ vh).transpose(1, 0, 2).reshape(S, D)
        np.testing.assert_allclose(out[b], ref, rtol=1e-4, atol=1e-5)


def test_decode_attention_self_and_cross(simt):
    rs = np.random.RandomState(4)
    B, H, S, max_t = 3, 2, 333, 32
    D = H * 64
    kc = np.zeros((B, H, max_t, 64), np.float32)
    vc = np.zeros_like(kc)
    hist = []
    for pos in range(5):
        qkv = rs.randn(B, 3 * D).astype(np.float32)
        hist.append(qkv)
        out = np.zeros((B, D), np.float32)
        simt.emu_decode_self(P(qkv), P(kc), P(vc), P(out), pos, max_t, H, B)
        K = np.stack([h[:, D:2 * D] for h in hist], 1).reshape(B, pos + 1, H, 64)
        V = np.stack([h[:, 2 * D:] for h in hist], 1).reshape(B, pos + 1, H, 64)
        q = qkv[:, :D].reshape(B, H, 64) * 0.125
        s = np.einsum("bhe,bthe->bht", q, K)
        ref = np.einsum("bht,bthe->bhe", model_ref.softmax(s), V).reshape(B, D)
        np.testing.assert_allclose(out, ref, rtol=1e-4, atol=1e-5)
    # cross: [B*S][2D] -> per-head streams -> split-S attention with ticketed merge
    kv = rs.randn(B * S, 2 * D).astype(np.float32)
    kx = np.zeros((B, H, S, 64), np.float32)
    vx = np.zeros_like(kx)
    simt.emu_split_heads(P(kv), P(kx), P(vx), B, S, H)
    np.testing.assert_array_equal(kx, kv[:, :D].reshape(B, S, H, 64).transpose(0, 2, 1, 3))
    q = rs.randn(B, D).astype(np.float32)
    part = np.zeros((B * H * 16 * 66,), np.float32)
    tickets = np.zeros((B * H,), np.uint32)
    for nsplit in (1, 5, 16):
        out = np.zeros((B, D), np.float32)
        simt.emu_decode_cross(P(q), P(kx), P(vx), P(out), P(part), tickets.ctypes.data_as(ctypes.c_void_p), S, H, B,
                              nsplit)
        s = np.einsum("bhe,bhse->bhs", q.reshape(B, H, 64) * 0.125, kx)
        ref = np.einsum("bhs,bhse->bhe", model_ref.softmax(s), vx).reshape(B, D)
        np.testing.assert_allclose(out, ref, rtol=1e-4, atol=1e-5)
        assert np.all(tickets == 0)


def test_argmax_last_tie_and_eot_state(simt):
    """whisper.cpp:346-361: `>=` scan from max_value = row[0]: last maximal index; a NaN at index > 0 never wins, a NaN
    at index 0 is never replaced (the oracle's argmax_last restates the same scan)."""
    B, V = 4, 1000
    logits = np.random.RandomState(5).randn(B, V).astype(np.float32)
    logits[0, 10] = logits[0, 900] = 9.0  # tie -> last index
    logits[1, 1] = np.nan
    logits[1, 77] = 5.0
    logits[2, 500] = 8.0
    logits[3, 0] = np.nan                 # NaN first: index 0 survives every comparison
    logits[3, 400] = 9.0
    tokens = np.full((B, 31), -1, np.int32)
    lens = np.full(B, 4, np.int32)
    done = np.array([0, 0, 1, 0], np.int32)
    I = lambda a: a.ctypes.data_as(ip)
    simt.emu_argmax_append(P(logits), V, I(tokens), 31, I(lens), I(done), 3, 77, B)
    assert tokens[0, 4] == 900 and lens[0] == 5 and done[0] == 0
    assert tokens[1, 4] == 77 and done[1] == 1  # NaN at index 1 never wins; 77 == eot -> done
    assert tokens[2, 4] == -1 and lens[2] == 4  # finished clip untouched
    assert tokens[3, 4] == 0 and lens[3] == 5
    for b in (0, 1, 3):
        assert model_ref.argmax_last(logits[b]) == tokens[b, 4]


def test_mel_to_time_major(simt):
    rs = np.random.RandomState(6)
    B, n_mel, n_len = 2, 80, 70
    mel = rs.randn(B, n_mel, n_len).astype(np.float32)
    out = np.zeros((B, n_len + 2, n_mel), np.float32)
    simt.emu_mel_to_tm(P(mel), None, P(out), n_mel, n_len, B)
    np.testing.assert_array_equal(out[:, 1:-1], mel.transpose(0, 2, 1))
    assert np.all(out[:, 0] == 0) and np.all(out[:, -1] == 0)
