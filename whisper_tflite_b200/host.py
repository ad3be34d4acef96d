"""ctypes access to csrc/host/libwhisper_host.so — the C++ mirror of the reference's whisper.h
(Engine / EncDec / Monolith / create_engine / Reader / decode / log_mel_spectrogram / wav_read_legacy /
MmapFile) through the flat `wth_*` shim in csrc/host/host_capi.cpp. Tests and tools drive the host
layer through this exactly as app/encdec or the JNI bridge would; nothing here computes anything."""
from __future__ import annotations

import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
HOST_DIR = os.path.join(HERE, "csrc", "host")
HOST_LIB = os.path.join(HOST_DIR, "libwhisper_host.so")
ENCDEC_EXE = os.path.join(HOST_DIR, "encdec")
MINIMAL_EXE = os.path.join(HOST_DIR, "minimal")

ENGINE_MONOLITH, ENGINE_ENCDEC = 0, 1  # whisper.h:199-204
SPECIAL_KEYS = ("eot", "sot", "translate", "transcribe", "prev", "solm", "not", "beg")

EXPORTED = [
    "wth_last_error", "wth_blob_read", "wth_blob_free", "wth_blob_info", "wth_blob_filters", "wth_blob_token",
    "wth_blob_decode", "wth_blob_decode_i32", "wth_language_id", "wth_lang_code", "wth_remove_extra_spaces",
    "wth_wav_read_legacy", "wth_wav_read", "wth_mmap_probe", "wth_log_mel", "wth_engine_create",
    "wth_engine_destroy", "wth_engine_transcribe", "wth_engine_transcribe_file", "wth_engine_transcribe_batch",
]

_lib = None
_f32p = ctypes.POINTER(ctypes.c_float)


def load() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(HOST_LIB):
            raise RuntimeError(f"{HOST_LIB} is missing: run `python -m whisper_tflite_b200.build`")
        lib = ctypes.CDLL(HOST_LIB)
        lib.wth_last_error.restype = ctypes.c_char_p
        lib.wth_blob_read.restype = ctypes.c_void_p
        lib.wth_blob_read.argtypes = [ctypes.c_void_p, ctypes.c_int]
        lib.wth_engine_create.restype = ctypes.c_void_p
        lib.wth_engine_create.argtypes = [ctypes.c_int, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int]
        lib.wth_mmap_probe.restype = ctypes.c_longlong
        for fn in ("wth_blob_free", "wth_engine_destroy"):
            getattr(lib, fn).restype = None
            getattr(lib, fn).argtypes = [ctypes.c_void_p]
        _lib = lib
    return _lib


def _s(fn, *args, cap=1 << 16) -> bytes:
    buf = ctypes.create_string_buffer(cap)
    n = fn(*args, buf, cap)
    if n < 0:
        raise RuntimeError(f"host call failed ({n})")
    return buf.raw[:n]


class Blob:
    """whisper::Reader on a vocab+filters blob (bytes include the u64 size prefix, skipped like whisper.cpp:676-681)."""

    def __init__(self, blob: bytes, multilingual: bool):
        self.lib = load()
        self._buf = ctypes.create_string_buffer(blob, len(blob))
        self.h = ctypes.c_void_p(self.lib.wth_blob_read(ctypes.addressof(self._buf) + 8, int(multilingual)))
        nv, nm, nf = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        sp = (ctypes.c_int * 8)()
        self.n_entries = self.lib.wth_blob_info(self.h, ctypes.byref(nv), ctypes.byref(nm), ctypes.byref(nf), sp)
        self.n_vocab, self.n_mel, self.n_fft = nv.value, nm.value, nf.value
        self.special = dict(zip(SPECIAL_KEYS, list(sp)))
        f = np.empty(self.n_mel * self.n_fft, dtype=np.float32)
        self.lib.wth_blob_filters(self.h, f.ctypes.data_as(_f32p))
        self.filters = f.reshape(self.n_mel, self.n_fft)

    def token(self, i: int):
        buf = ctypes.create_string_buffer(512)
        n = self.lib.wth_blob_token(self.h, int(i), buf, 512)
        return None if n < 0 else buf.raw[:n]

    def decode(self, ids, omit_special: bool = False, i32: bool = False) -> bytes:
        if i32:
            arr = (ctypes.c_int32 * len(ids))(*[int(i) for i in ids])
            return _s(self.lib.wth_blob_decode_i32, self.h, arr, len(ids), int(omit_special))
        arr = (ctypes.c_int64 * len(ids))(*[int(i) for i in ids])
        return _s(self.lib.wth_blob_decode, self.h, arr, len(ids), int(omit_special))

    def close(self):
        if self.h:
            self.lib.wth_blob_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def language_id(code: str) -> int:
    return int(load().wth_language_id(code.encode()))


def lang_code(i: int) -> str:
    return _s(load().wth_lang_code, int(i), cap=64).decode()


def remove_extra_spaces(s: str) -> str:
    return _s(load().wth_remove_extra_spaces, s.encode(), cap=len(s.encode()) + 8).decode()


def wav_read_legacy(path: str, cap: int = 1 << 24) -> np.ndarray:
    out = np.zeros(cap, dtype=np.float32)
    n = load().wth_wav_read_legacy(path.encode(), out.ctypes.data_as(_f32p), cap)
    return out[:min(n, cap)].copy()


def wav_read(path: str, cap: int = 1 << 24) -> np.ndarray:
    out = np.zeros(cap, dtype=np.float32)
    n = load().wth_wav_read(path.encode(), out.ctypes.data_as(_f32p), cap)
    return out[:min(n, cap)].copy()


def mmap_probe(path: str):
    """-> (size, first 16 bytes) through whisper::MmapFile, or raises RuntimeError with its message."""
    first = (ctypes.c_ubyte * 16)()
    n = load().wth_mmap_probe(path.encode(), first)
    if n < 0:
        raise RuntimeError(load().wth_last_error().decode())
    return int(n), bytes(first)[:min(16, n)]


def log_mel_spectrogram(pcm: np.ndarray, filters: np.ndarray) -> np.ndarray:
    """whisper::log_mel_spectrogram (device-computed) for one clip -> [n_mel][n_samples/160] f32."""
    pcm = np.ascontiguousarray(pcm, dtype=np.float32)
    filters = np.ascontiguousarray(filters, dtype=np.float32)
    out = np.empty((filters.shape[0], pcm.shape[0] // 160), dtype=np.float32)
    n_len = ctypes.c_int()
    rc = load().wth_log_mel(pcm.ctypes.data_as(_f32p), pcm.shape[0], filters.ctypes.data_as(_f32p), filters.shape[0],
                            filters.shape[1], out.ctypes.data_as(_f32p), ctypes.byref(n_len))
    if rc != 0:
        raise RuntimeError("log_mel_spectrogram returned false")
    assert n_len.value == out.shape[1]
    return out


class HostEngine:
    """whisper::create_engine(...) + Engine::transcribe / transcribe_batch."""

    def __init__(self, engine_type: int, model_prefix: str, vocab_path: str, multilingual: bool):
        self.lib = load()
        h = self.lib.wth_engine_create(int(engine_type), model_prefix.encode(), vocab_path.encode(), int(multilingual))
        if not h:
            raise RuntimeError("create_engine returned nullptr: " + self.lib.wth_last_error().decode())
        self.h = ctypes.c_void_p(h)

    def transcribe(self, pcm: np.ndarray):
        """-> (text bytes, size of the caller's vector after the call: the reference resizes it to 480000)."""
        pcm = np.ascontiguousarray(pcm, dtype=np.float32)
        after = ctypes.c_int()
        buf = ctypes.create_string_buffer(1 << 16)
        n = self.lib.wth_engine_transcribe(self.h, pcm.ctypes.data_as(_f32p), pcm.shape[0], buf, 1 << 16, ctypes.byref(after))
        assert n >= 0
        return buf.raw[:n], after.value

    def transcribe_file(self, path: str) -> bytes:
        return _s(self.lib.wth_engine_transcribe_file, self.h, path.encode())

    def transcribe_batch(self, pcm: np.ndarray):
        pcm = np.ascontiguousarray(pcm, dtype=np.float32)
        cap = 1 << 20
        buf = ctypes.create_string_buffer(cap)
        n = self.lib.wth_engine_transcribe_batch(self.h, pcm.ctypes.data_as(_f32p), pcm.shape[0], pcm.shape[1], buf, cap)
        assert n >= 0
        parts = buf.raw[:n].split(b"\0")
        return parts[:pcm.shape[0]]

    def transcribe_batch_i16(self, pcm16: np.ndarray):
        """16-bit samples [n_clips][n_samples] (converted on the device: wav_util.cpp:70-73's `/ INT16_MAX`)."""
        pcm16 = np.ascontiguousarray(pcm16, dtype=np.int16)
        cap = 1 << 20
        buf = ctypes.create_string_buffer(cap)
        n = self.lib.wth_engine_transcribe_batch_i16(self.h, pcm16.ctypes.data_as(ctypes.c_void_p), pcm16.shape[0],
                                                     pcm16.shape[1], buf, cap)
        assert n >= 0
        return buf.raw[:n].split(b"\0")[:pcm16.shape[0]]

    def set_language(self, code: str) -> bool:
        """Language tag of the prompt (reference: hard-coded "de", whisper.cpp:327). False for an unknown code."""
        return bool(self.lib.wth_engine_set_language(self.h, code.encode()))

    def set_task(self, translate: bool):
        self.lib.wth_engine_set_task(self.h, int(translate))

    def set_omit_special_tokens(self, omit: bool):
        self.lib.wth_engine_set_omit_special_tokens(self.h, int(omit))

    def close(self):
        if self.h:
            self.lib.wth_engine_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
