"""Synthetic assets for the Whisper hot path: PCM clips, the slaney mel filterbank, seeded
random-init weights of the named architecture, the engine weight file (``<prefix>.b200w``)
and the reference's ``filters_vocab_*.bin`` blob.

Nothing ships with the reference (no weights, vocab blob or audio: scripts/run.sh:7-16), and
there is no network, so the benchmark and the parity tests use the fixtures BASELINE.md §3
defines: seed 0; per clip ``0.3*sin(2*pi*f_c*t) + 0.1*N(0,1)``, ``f_c ~ U[100, 4000]`` Hz,
trailing 0-5 s zeroed, 480 000 f32 samples at 16 kHz.

This module is pure numpy and has no dependency on ``oracle/``.
"""
from __future__ import annotations

import dataclasses
import struct
from typing import Dict

import numpy as np

SAMPLE_RATE = 16000  # whisper.h:34
N_FFT = 400  # whisper.h:35
N_MEL = 80  # whisper.h:36
HOP = 160  # whisper.h:37
CHUNK_S = 30  # whisper.h:38
N_FRAMES = 3000  # whisper.h:39
N_SAMPLES = SAMPLE_RATE * CHUNK_S
N_FREQ = 1 + N_FFT // 2  # whisper.cpp:129

VOCAB_EN = 51864  # whisper.h:41
VOCAB_MULTI = 51865  # whisper.h:42


@dataclasses.dataclass(frozen=True)
class WhisperDims:
    """Architecture dims. head_dim is 64 everywhere."""

    name: str
    d_model: int
    n_heads: int
    n_enc_layers: int
    n_dec_layers: int
    n_vocab: int = VOCAB_MULTI
    n_mels: int = N_MEL
    n_audio_ctx: int = 1500
    n_text_ctx: int = 448

    @property
    def head_dim(self) -> int:
        return self.d_model // self.n_heads


DIMS = {
    # "micro" is not a Whisper size: it is the smallest shape the kernels accept (d multiple of
    # 64, head_dim 64) and keeps CPU-oracle tests to a fraction of a second.
    "micro": WhisperDims("micro", 128, 2, 2, 2),
    "tiny": WhisperDims("tiny", 384, 6, 4, 4),
    "base": WhisperDims("base", 512, 8, 6, 6),
    "small": WhisperDims("small", 768, 12, 12, 12),
}


# --------------------------------------------------------------------------------------
# audio
# --------------------------------------------------------------------------------------
def synth_pcm(n_clips: int, seed: int = 0, n_samples: int = N_SAMPLES) -> np.ndarray:
    """[n_clips, n_samples] f32 in [-1, 1]. Clip i depends only on (seed, i)."""
    out = np.empty((n_clips, n_samples), dtype=np.float32)
    t = np.arange(n_samples, dtype=np.float64) / SAMPLE_RATE
    for i in range(n_clips):
        rs = np.random.RandomState((seed * 1000003 + i) & 0x7FFFFFFF)
        f_c = rs.uniform(100.0, 4000.0)
        tail_s = rs.uniform(0.0, 5.0)
        x = 0.3 * np.sin(2.0 * np.pi * f_c * t) + 0.1 * rs.standard_normal(n_samples)
        x = np.clip(x, -1.0, 1.0)
        n_zero = int(tail_s * SAMPLE_RATE)
        if n_zero > 0:
            x[n_samples - n_zero:] = 0.0
        out[i] = x.astype(np.float32)
    return out


def _hz_to_mel(f):
    f = np.asarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-10) / min_log_hz) / logstep, f / f_sp)


def _mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), m * f_sp)


def mel_filters(n_mel: int = N_MEL, n_freq: int = N_FREQ, sr: int = SAMPLE_RATE) -> np.ndarray:
    """Slaney-scale, slaney-normalised triangular filterbank, ``[n_mel][n_freq]`` f32 row-major —
    the layout of ``Filters.data`` (whisper.h:96-101, indexed ``j*n_fft+k`` at whisper.cpp:173).
    Bit-identical in f32 to HF ``WhisperFeatureExtractor().mel_filters.T`` and to OpenAI's
    ``mel_filters.npz`` (checked in tests/test_synth.py)."""
    fft_freqs = np.linspace(0, sr / 2, n_freq)
    pts = _mel_to_hz(np.linspace(_hz_to_mel(0.0), _hz_to_mel(sr / 2), n_mel + 2))
    fdiff = np.diff(pts)
    slopes = pts[None, :] - fft_freqs[:, None]
    down = -slopes[:, :-2] / fdiff[:-1]
    up = slopes[:, 2:] / fdiff[1:]
    w = np.maximum(0, np.minimum(down, up))
    w = w * (2.0 / (pts[2:n_mel + 2] - pts[:n_mel]))[None, :]
    return np.ascontiguousarray(w.T.astype(np.float32))


# --------------------------------------------------------------------------------------
# weights
# --------------------------------------------------------------------------------------
def sinusoids(length: int, channels: int, max_timescale: float = 10000.0) -> np.ndarray:
    """Encoder positional table (openai-whisper ``sinusoids``; HF modeling_whisper.py:55)."""
    inc = np.log(max_timescale) / (channels // 2 - 1)
    inv = np.exp(-inc * np.arange(channels // 2, dtype=np.float64))
    t = np.arange(length, dtype=np.float64)[:, None] * inv[None, :]
    return np.concatenate([np.sin(t), np.cos(t)], axis=1).astype(np.float32)


def weight_names(dims: WhisperDims):
    """Canonical tensor names and shapes (openai-whisper naming: the reference's graphs are exported
    from ``whisper.load_model("tiny").encoder/.decoder`` — export/generate_onnx.py:85-130)."""
    d, V = dims.d_model, dims.n_vocab
    out = [
        ("encoder.conv1.weight", (d, dims.n_mels, 3)), ("encoder.conv1.bias", (d,)),
        ("encoder.conv2.weight", (d, d, 3)), ("encoder.conv2.bias", (d,)),
        ("encoder.positional_embedding", (dims.n_audio_ctx, d)),
    ]

    def attn(p):
        return [(p + ".query.weight", (d, d)), (p + ".query.bias", (d,)),
                (p + ".key.weight", (d, d)),  # key has no bias
                (p + ".value.weight", (d, d)), (p + ".value.bias", (d,)),
                (p + ".out.weight", (d, d)), (p + ".out.bias", (d,))]

    def ln(p):
        return [(p + ".weight", (d,)), (p + ".bias", (d,))]

    def mlp(p):
        return [(p + ".0.weight", (4 * d, d)), (p + ".0.bias", (4 * d,)),
                (p + ".2.weight", (d, 4 * d)), (p + ".2.bias", (d,))]

    for i in range(dims.n_enc_layers):
        p = f"encoder.blocks.{i}"
        out += ln(p + ".attn_ln") + attn(p + ".attn") + ln(p + ".mlp_ln") + mlp(p + ".mlp")
    out += ln("encoder.ln_post")
    out += [("decoder.token_embedding.weight", (V, d)), ("decoder.positional_embedding", (dims.n_text_ctx, d))]
    for i in range(dims.n_dec_layers):
        p = f"decoder.blocks.{i}"
        out += (ln(p + ".attn_ln") + attn(p + ".attn") + ln(p + ".cross_attn_ln") + attn(p + ".cross_attn")
                + ln(p + ".mlp_ln") + mlp(p + ".mlp"))
    out += ln("decoder.ln")
    return out


def init_weights(dims: WhisperDims, seed: int = 0) -> Dict[str, np.ndarray]:
    """Seeded random init of the named architecture, f32. Fan-in scaled (std = 1/sqrt(fan_in)) so the
    signal reaches the logits with O(1) spread and the greedy margin is not degenerate
    (SURVEY.md §7.2 item 4); LN gains near 1; encoder positions are the fixed sinusoids."""
    rs = np.random.RandomState(seed)
    w: Dict[str, np.ndarray] = {}
    for name, shape in weight_names(dims):
        if name == "encoder.positional_embedding":
            a = sinusoids(*shape)
        elif name == "decoder.token_embedding.weight":
            a = rs.standard_normal(shape).astype(np.float32) * np.float32(0.08)
        elif name == "decoder.positional_embedding":
            a = rs.standard_normal(shape).astype(np.float32) * np.float32(0.04)
        elif name.endswith("_ln.weight") or name.endswith("ln_post.weight") or name == "decoder.ln.weight":
            a = (1.0 + 0.1 * rs.standard_normal(shape)).astype(np.float32)
        elif name.endswith(".bias"):
            a = (0.05 * rs.standard_normal(shape)).astype(np.float32)
        else:
            fan_in = int(np.prod(shape[1:]))
            a = (rs.standard_normal(shape) / np.sqrt(fan_in)).astype(np.float32)
        w[name] = np.ascontiguousarray(a)
    return w


# --------------------------------------------------------------------------------------
# engine weight file  <prefix>.b200w
# --------------------------------------------------------------------------------------
WEIGHT_MAGIC = b"WTB200W\x01"


def write_weight_file(path: str, dims: WhisperDims, weights: Dict[str, np.ndarray]) -> None:
    """Flat little-endian container read by ``wt_engine_create`` (csrc/weights.cpp):
    magic[8] | u32 version | u32 n_tensors | i32 dims[8] | index | 256-aligned f32 payloads.
    Index entry: u16 name_len | name | u8 dtype(0=f32) | u8 ndim | i64 shape[ndim] | u64 offset | u64 nbytes."""
    names = [n for n, _ in weight_names(dims)]
    index = bytearray()
    entries = []
    for n in names:
        a = np.ascontiguousarray(weights[n], dtype=np.float32)
        entries.append((n, a))
    header_fixed = len(WEIGHT_MAGIC) + 4 + 4 + 8 * 4
    idx_size = sum(2 + len(n.encode()) + 1 + 1 + 8 * a.ndim + 8 + 8 for n, a in entries)
    off = (header_fixed + idx_size + 255) // 256 * 256
    offsets = []
    for n, a in entries:
        offsets.append(off)
        off = (off + a.nbytes + 255) // 256 * 256
    for (n, a), o in zip(entries, offsets):
        nb = n.encode()
        index += struct.pack("<H", len(nb)) + nb + struct.pack("<BB", 0, a.ndim)
        index += struct.pack("<%dq" % a.ndim, *a.shape) + struct.pack("<QQ", o, a.nbytes)
    assert len(index) == idx_size
    with open(path, "wb") as f:
        f.write(WEIGHT_MAGIC)
        f.write(struct.pack("<II", 1, len(entries)))
        f.write(struct.pack("<8i", dims.n_mels, dims.n_audio_ctx, dims.d_model, dims.n_heads,
                            dims.n_enc_layers, dims.n_dec_layers, dims.n_vocab, dims.n_text_ctx))
        f.write(bytes(index))
        for (n, a), o in zip(entries, offsets):
            f.seek(o)
            f.write(a.tobytes())
        f.truncate(off)


def read_weight_file(path: str):
    """Inverse of :func:`write_weight_file` (used by tests and tools; the engine has its own C++ reader)."""
    with open(path, "rb") as f:
        blob = f.read()
    assert blob[:8] == WEIGHT_MAGIC, "bad magic"
    version, n = struct.unpack_from("<II", blob, 8)
    d = struct.unpack_from("<8i", blob, 16)
    dims = WhisperDims("file", d_model=d[2], n_heads=d[3], n_enc_layers=d[4], n_dec_layers=d[5], n_vocab=d[6],
                       n_mels=d[0], n_audio_ctx=d[1], n_text_ctx=d[7])
    p = 48
    out = {}
    for _ in range(n):
        (ln,) = struct.unpack_from("<H", blob, p); p += 2
        name = blob[p:p + ln].decode(); p += ln
        dt, nd = struct.unpack_from("<BB", blob, p); p += 2
        shape = struct.unpack_from("<%dq" % nd, blob, p); p += 8 * nd
        off, nb = struct.unpack_from("<QQ", blob, p); p += 16
        out[name] = np.frombuffer(blob, dtype=np.float32, count=nb // 4, offset=off).reshape(shape)
    return dims, out


# --------------------------------------------------------------------------------------
# vocab + filters blob (whisper.h:214-235, whisper.cpp:519-575, :676-681)
# --------------------------------------------------------------------------------------
def synth_vocab_tokens(n_tokens: int = 50257):
    """Stand-in token strings for ids < n_tokens (the real GPT-2 byte-pair table is not on this box).
    Token i decodes to a short ASCII word with a leading space so decoded text is readable."""
    return [(" t%d" % i).encode() for i in range(n_tokens)]


def write_vocab_blob(path: str, filters: np.ndarray, tokens, magic: int = 0x74666C74) -> int:
    """``u64 size | u32 magic | i32 n_mel | i32 n_fft | f32[n_mel*n_fft] | i32 n_vocab | {u32 len, bytes}*``.
    Returns the file size. The u64 prefix holds the payload size and is skipped by the engines."""
    n_mel, n_fft = filters.shape
    body = bytearray()
    body += struct.pack("<Iii", magic, n_mel, n_fft)
    body += np.ascontiguousarray(filters, dtype=np.float32).tobytes()
    body += struct.pack("<i", len(tokens))
    for t in tokens:
        assert len(t) <= 255  # whisper.cpp:563-571 reads into a 256-byte buffer
        body += struct.pack("<I", len(t)) + t
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(body)))
        f.write(bytes(body))
    return 8 + len(body)
