"""ctypes binding of include/whisper_b200.h — the call a Python user makes.

Host arrays are numpy; device arrays are anything exposing ``data_ptr()`` (torch CUDA tensors).
There is no CPU implementation behind these calls: if the CUDA library is missing or no B200 is
visible, construction raises.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional, Sequence

import numpy as np

from . import synth

HERE = os.path.dirname(os.path.abspath(__file__))
# WT_LIB selects another in-tree build of the same sources (e.g. the -DDS_TRACE_FINE=1 variant tools/dec_trace.py builds)
LIB_PATH = os.environ.get("WT_LIB") or os.path.join(HERE, "csrc", "libwhisper_b200.so")

WT_HOST, WT_DEVICE = 0, 1
PRECISION_BF16, PRECISION_FP32_VERIFY = 0, 1
REFERENCE_PROMPT = (50258, 50261, 50359, 50363)  # whisper.cpp:331 (multilingual, "de")
EOT_MULTILINGUAL = 50257
MAX_TOKENS = 31

EXPORTED = [
    "wt_engine_create", "wt_engine_destroy", "wt_engine_dims", "wt_set_mel_filters", "wt_set_stream",
    "wt_synchronize", "wt_log_mel", "wt_encode", "wt_decode_greedy", "wt_transcribe_tokens",
    "wt_get_step_logits", "wt_launch_count", "wt_last_error", "wt_version", "wt_debug_gemm_bf16",
    "wt_profile_enable", "wt_profile_read", "wt_stage_count", "wt_stage_name", "wt_debug_decoder_trace",
    "wt_set_decoder_group", "wt_debug_hash", "wt_debug_gemm_ex", "wt_debug_attention_bf16", "wt_log_mel_i16",
    "wt_transcribe_tokens_i16", "wt_gemm_argmax_noalloc",
]


class WtConfig(ctypes.Structure):
    _fields_ = [("struct_size", ctypes.c_int32), ("precision", ctypes.c_int32), ("max_batch", ctypes.c_int32),
                ("keep_logits", ctypes.c_int32), ("decoder_cluster", ctypes.c_int32)]


class WtDims(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in ("n_mels", "n_audio_ctx", "d_model", "n_heads", "n_enc_layers",
                                              "n_dec_layers", "n_vocab", "n_text_ctx")]


class WhisperB200Error(RuntimeError):
    pass


_lib = None


def load_library() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise WhisperB200Error(
                f"{LIB_PATH} is missing: build it with `python -m whisper_tflite_b200.build` "
                "(the CUDA library is the product; there is no CPU fallback)")
        lib = ctypes.CDLL(LIB_PATH)
        lib.wt_last_error.restype = ctypes.c_char_p
        lib.wt_version.restype = ctypes.c_char_p
        lib.wt_launch_count.restype = ctypes.c_int64
        lib.wt_launch_count.argtypes = [ctypes.c_void_p]
        lib.wt_stage_name.restype = ctypes.c_char_p
        lib.wt_engine_destroy.restype = None
        lib.wt_engine_destroy.argtypes = [ctypes.c_void_p]
        _lib = lib
    return _lib


def _check(rc: int) -> None:
    if rc != 0:
        raise WhisperB200Error(f"wt_status {rc}: {load_library().wt_last_error().decode()}")


def _ptr_mem(a):
    """-> (address, WT_HOST|WT_DEVICE, keepalive)"""
    if a is None:
        return None, WT_DEVICE, None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data, WT_HOST, a
    if hasattr(a, "data_ptr"):  # torch tensor
        assert a.is_contiguous()
        return a.data_ptr(), (WT_DEVICE if a.is_cuda else WT_HOST), a
    raise TypeError(type(a))


def _is_i16(a) -> bool:
    return str(getattr(a, "dtype", "")).replace("torch.", "") == "int16"


class Engine:
    """One engine per device; not re-entrant (same contract as the reference's Engine)."""

    def __init__(self, weights_path: Optional[str], device: int = 0, max_batch: int = 1,
                 precision: int = PRECISION_BF16, keep_logits: bool = False, filters: Optional[np.ndarray] = None,
                 decoder_cluster: int = 0):
        self.lib = load_library()
        cfg = WtConfig(ctypes.sizeof(WtConfig), precision, max_batch, int(keep_logits), int(decoder_cluster))
        h = ctypes.c_void_p()
        _check(self.lib.wt_engine_create(ctypes.byref(cfg), weights_path.encode() if weights_path else None,
                                         int(device), ctypes.byref(h)))
        self.h = h
        self.max_batch = max_batch
        self.precision = precision
        self.device = device
        self.dims = None
        if weights_path:
            d = WtDims()
            _check(self.lib.wt_engine_dims(self.h, ctypes.byref(d)))
            self.dims = d
        self.set_mel_filters(synth.mel_filters() if filters is None else filters)

    def close(self):
        if getattr(self, "h", None):
            self.lib.wt_engine_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_mel_filters(self, filters: np.ndarray):
        f = np.ascontiguousarray(filters, dtype=np.float32)
        _check(self.lib.wt_set_mel_filters(self.h, ctypes.c_void_p(f.ctypes.data), f.shape[0], f.shape[1]))

    def set_stream(self, cuda_stream: int):
        _check(self.lib.wt_set_stream(self.h, ctypes.c_void_p(cuda_stream)))

    def set_decoder_group(self, clips_per_cluster: int):
        """Clips per decoder cluster (0 = automatic, 16 = fewest SMs per batch: for pipelined serving)."""
        _check(self.lib.wt_set_decoder_group(self.h, int(clips_per_cluster)))

    def synchronize(self):
        _check(self.lib.wt_synchronize(self.h))

    @property
    def launch_count(self) -> int:
        return int(self.lib.wt_launch_count(self.h))

    # ---- whisper::log_mel_spectrogram, batched ----
    def log_mel(self, pcm, out=None):
        p, mem, _k = _ptr_mem(pcm)
        n_clips, n_samples = pcm.shape
        if out is None:
            out = np.empty((n_clips, 80, n_samples // 160), dtype=np.float32)
        o, omem, _k2 = _ptr_mem(out)
        fn = self.lib.wt_log_mel_i16 if _is_i16(pcm) else self.lib.wt_log_mel   # int16 samples: converted on the device
        _check(fn(self.h, ctypes.c_void_p(p), n_clips, n_samples, mem, ctypes.c_void_p(o), omem))
        return out

    # ---- Encoder::forward ----
    def encode(self, mel=None, n_clips: Optional[int] = None, out=None, want_states: bool = True):
        p, mem, _k = _ptr_mem(mel)
        if mel is not None:
            n_clips = mel.shape[0]
        if out is None and want_states and self.dims is not None:
            out = np.empty((n_clips, self.dims.n_audio_ctx, self.dims.d_model), dtype=np.float32)
        o, omem, _k2 = _ptr_mem(out)
        _check(self.lib.wt_encode(self.h, ctypes.c_void_p(p) if p else None, n_clips, mem,
                                  ctypes.c_void_p(o) if o else None, omem))
        return out

    # ---- Decoder::forward ----
    def decode_greedy(self, n_clips: int, prompt: Sequence[int] = REFERENCE_PROMPT, max_total_tokens: int = MAX_TOKENS,
                      eot: int = EOT_MULTILINGUAL, ids_out=None, lens_out=None):
        pr = (ctypes.c_int64 * len(prompt))(*[int(t) for t in prompt])
        if ids_out is None:
            ids_out = np.empty((n_clips, max_total_tokens), dtype=np.int32)
            lens_out = np.empty((n_clips,), dtype=np.int32)
        i, imem, _k = _ptr_mem(ids_out)
        l, lmem, _k2 = _ptr_mem(lens_out)
        assert imem == lmem
        _check(self.lib.wt_decode_greedy(self.h, n_clips, pr, len(prompt), max_total_tokens, ctypes.c_int64(eot),
                                         ctypes.c_void_p(i), ctypes.c_void_p(l), imem))
        return ids_out, lens_out

    # ---- Monolith / EncDec transcribe minus detokenisation, batched ----
    def transcribe_tokens(self, pcm, prompt: Sequence[int] = REFERENCE_PROMPT, max_total_tokens: int = MAX_TOKENS,
                          eot: int = EOT_MULTILINGUAL, ids_out=None, lens_out=None):
        p, mem, _k = _ptr_mem(pcm)
        n_clips, n_samples = pcm.shape
        pr = (ctypes.c_int64 * len(prompt))(*[int(t) for t in prompt])
        if ids_out is None:
            ids_out = np.empty((n_clips, max_total_tokens), dtype=np.int32)
            lens_out = np.empty((n_clips,), dtype=np.int32)
        i, imem, _k2 = _ptr_mem(ids_out)
        l, lmem, _k3 = _ptr_mem(lens_out)
        assert imem == lmem
        fn = self.lib.wt_transcribe_tokens_i16 if _is_i16(pcm) else self.lib.wt_transcribe_tokens
        _check(fn(self.h, ctypes.c_void_p(p), n_clips, n_samples, mem, pr, len(prompt), max_total_tokens,
                  ctypes.c_int64(eot), ctypes.c_void_p(i), ctypes.c_void_p(l), imem))
        return ids_out, lens_out

    def profile(self, enable: bool):
        _check(self.lib.wt_profile_enable(self.h, int(enable)))

    def profile_read(self):
        """-> {stage name: (total ms, launches)} accumulated since profile(True)."""
        n = self.lib.wt_stage_count()
        ms = (ctypes.c_float * n)()
        cnt = (ctypes.c_int32 * n)()
        _check(self.lib.wt_profile_read(self.h, ms, cnt, n))
        return {self.lib.wt_stage_name(i).decode(): (float(ms[i]), int(cnt[i])) for i in range(n)}

    def decoder_trace(self, reset: bool = True):
        """WT_DEC_TRACE=1 engines: {phase: (total us, barrier-wait us)} of the persistent decoder as seen by CTA 0."""
        names = ["init", "ln1_qkv", "self_attn", "out_proj", "lnx_xq", "cross_attn", "xout_proj", "ln2_fc1", "fc2",
                 "logits", "finalize_embed", "gemm:norm", "gemm:mma", "gemm:epilogue", "sync:fence", "sync:bar+arrive"]
        buf = (ctypes.c_uint64 * 32)()
        _check(self.lib.wt_debug_decoder_trace(self.h, buf, int(reset)))
        return {n: (buf[i] / 1e3, buf[16 + i] / 1e3) for i, n in enumerate(names)}

    def debug_hash(self, what: str, n_clips: int) -> int:
        """64-bit digest of a resident buffer ("enc_out", "cross_kv0", "x", "qkv", ... see whisper_b200.h)."""
        h = ctypes.c_uint64(0)
        _check(self.lib.wt_debug_hash(self.h, what.encode(), int(n_clips), ctypes.byref(h)))
        return int(h.value)

    def step_logits(self, step: int, n_clips: int) -> np.ndarray:
        out = np.empty((n_clips, self.dims.n_vocab), dtype=np.float32)
        _check(self.lib.wt_get_step_logits(self.h, step, n_clips, ctypes.c_void_p(out.ctypes.data), WT_HOST))
        return out


def debug_gemm_bf16(a, w, bias, c, flags: int = 0, stream: int = 0):
    """C[M][N] = epi(A[M][K] . W[N][K]^T) on torch CUDA bf16 tensors through the tcgen05 kernel."""
    lib = load_library()
    M, K = a.shape
    N = w.shape[0]
    _check(lib.wt_debug_gemm_bf16(ctypes.c_void_p(stream), ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(w.data_ptr()),
                                  ctypes.c_void_p(bias.data_ptr()) if bias is not None else None,
                                  ctypes.c_void_p(c.data_ptr()), M, N, K, flags | (1 if bias is not None else 0)))
    return c


EPI_BIAS, EPI_GELU, EPI_ADD, EPI_OUT_F32, EPI_ARGMAX, EPI_SPLIT_KV = 1, 2, 4, 16, 32, 64


class WtDebugGemmArgs(ctypes.Structure):
    _fields_ = [("struct_size", ctypes.c_int32),
                ("A", ctypes.c_void_p), ("a_dims", ctypes.c_uint64 * 3), ("a_strides", ctypes.c_uint64 * 2),
                ("rows_per_batch", ctypes.c_int32), ("n_batches", ctypes.c_int32),
                ("W", ctypes.c_void_p), ("bias", ctypes.c_void_p), ("N", ctypes.c_int32), ("K_total", ctypes.c_int32),
                ("taps", ctypes.c_int32), ("tap_k", ctypes.c_int32), ("tap_c0", ctypes.c_int32 * 3),
                ("tap_row", ctypes.c_int32 * 3),
                ("C", ctypes.c_void_p), ("c_batch_stride", ctypes.c_int64), ("c_rows", ctypes.c_int32),
                ("c_row_stride", ctypes.c_int32),
                ("addend", ctypes.c_void_p), ("add_batch_stride", ctypes.c_int64), ("add_rows", ctypes.c_int32),
                ("add_row_stride", ctypes.c_int32),
                ("argmax_part", ctypes.c_void_p),
                ("kv_out", ctypes.c_void_p), ("kv_out_v", ctypes.c_void_p), ("kv_S", ctypes.c_int32),
                ("kv_H", ctypes.c_int32), ("kv_chunked", ctypes.c_int32),
                ("flags", ctypes.c_int32), ("force_bn", ctypes.c_int32)]


def debug_gemm_ex(stream: int = 0, **kw):
    """wt_debug_gemm_ex with keyword fields of wt_debug_gemm_args; torch CUDA tensors are passed by data_ptr()."""
    a = WtDebugGemmArgs()
    a.struct_size = ctypes.sizeof(WtDebugGemmArgs)
    keep = []
    for k, v in kw.items():
        if hasattr(v, "data_ptr"):
            keep.append(v)
            v = v.data_ptr()
        if isinstance(v, (list, tuple)):
            arr = getattr(a, k)
            for i, x in enumerate(v):
                arr[i] = int(x)
        else:
            setattr(a, k, v)
    _check(load_library().wt_debug_gemm_ex(ctypes.c_void_p(stream), ctypes.byref(a)))


NOALLOC_MAX_PARTIALS = 131072   # wt_gemm_argmax_noalloc: M * ceil(N / bn) partial winners its static device scratch holds


def gemm_argmax_noalloc(a, w, bias, ids_out, stream: int = 0):
    """wt_gemm_argmax_noalloc: ids_out[m] = argmax_n (a[m] . w[n] + bias[n]) with the reference's tie / NaN rule, on torch
    CUDA tensors (a [M][K], w [N][K] bf16; bias [N] f32 or None; ids_out [M] int32). Allocates nothing on the device."""
    M, K = a.shape
    N = w.shape[0]
    _check(load_library().wt_gemm_argmax_noalloc(
        ctypes.c_void_p(stream), ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(w.data_ptr()),
        ctypes.c_void_p(bias.data_ptr()) if bias is not None else None, M, N, K, ctypes.c_void_p(ids_out.data_ptr())))
    return ids_out


def debug_attention_bf16(qkv, out, n_heads: int, stream: int = 0):
    """attention3_sm100_kernel on qkv [B][S][3*H*64] -> out [B][S][H*64] (torch CUDA bf16 tensors)."""
    B, S, _ = qkv.shape
    _check(load_library().wt_debug_attention_bf16(ctypes.c_void_p(stream), ctypes.c_void_p(qkv.data_ptr()),
                                                  ctypes.c_void_p(out.data_ptr()), B, S, n_heads))
    return out


class EnginePool:
    """Serving arrangement for one GPU: `n_engines` engines (replicated weights, one CUDA stream each) that take
    batches in turn, so batch i's greedy decoder — latency-bound, a few thread-block clusters — runs underneath batch
    i+1's mel frontend and encoder (tensor-bound, every SM the decoder leaves free). Each engine keeps the reference's
    contract (calls on ONE engine are serialised by its caller: Whisper.java:109-113); the pool only decides which
    engine a batch goes to. Host-buffer batches are driven by one thread per engine (the C call releases the GIL), so
    the H2D copy of a batch also overlaps the other engine's compute."""

    def __init__(self, weights_path: str, n_engines: int = 2, device: int = 0, max_batch: int = 64,
                 precision: int = PRECISION_BF16, decoder_group: int = 16, decoder_cluster: int = 0):
        from concurrent.futures import ThreadPoolExecutor
        self.engines = [Engine(weights_path, device=device, max_batch=max_batch, precision=precision,
                               decoder_cluster=decoder_cluster if n_engines > 1 else 0)
                        for _ in range(n_engines)]
        for e in self.engines:
            e.set_decoder_group(decoder_group if n_engines > 1 else 0)
        self._workers = [ThreadPoolExecutor(1) for _ in self.engines]
        self._next = 0

    def submit(self, pcm, ids_out=None, lens_out=None):
        """Queue one batch on the next engine; returns a future of (ids, lens)."""
        i = self._next
        self._next = (i + 1) % len(self.engines)
        return self._workers[i].submit(self.engines[i].transcribe_tokens, pcm, ids_out=ids_out, lens_out=lens_out)

    def transcribe_tokens_many(self, batches):
        """Token ids of every batch, in order."""
        return [f.result() for f in [self.submit(b) for b in batches]]

    @property
    def launch_count(self) -> int:
        return sum(e.launch_count for e in self.engines)

    def synchronize(self):
        for e in self.engines:
            e.synchronize()

    def close(self):
        for w in self._workers:
            w.shutdown(wait=True)
        for e in self.engines:
            e.close()
