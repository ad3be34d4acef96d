"""TEST INFRASTRUCTURE — ctypes access to the two CPU mel oracles.

* ``ref_*``    : oracle/_ref/libwhisper_ref.so = the reference's OWN sources
                 (/root/reference/whisper.tflite/whisper.cpp:37-216, :519-665; wav_util.cpp:18-87)
                 compiled in place by oracle/Makefile against the declaration-only TFLite stub.
* ``oracle_*`` : oracle/liboracle_mel.so = the plain-C restatement in oracle/mel_oracle.c.

Parity status: PINNED — the restatement is checked bit-for-bit against the compiled reference
(tests/test_oracle_mel.py) and both against tests/golden/mel_golden.npz.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libwhisper_ref.so")
ORACLE_SO = os.path.join(HERE, "liboracle_mel.so")

_f32p = ctypes.POINTER(ctypes.c_float)


def build(force: bool = False) -> None:
    """Compile the restatement, and the reference .so when /root/reference is present."""
    if force or not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(
            os.path.join(HERE, "mel_oracle.c")):
        subprocess.check_call(["make", "-s", "-C", HERE, "restatement"])
    if os.path.isdir("/root/reference/whisper.tflite") and (force or not os.path.exists(REF_SO)):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


def have_ref() -> bool:
    return os.path.exists(REF_SO)


_libs = {}


def _lib(path):
    if path not in _libs:
        _libs[path] = ctypes.CDLL(path)
    return _libs[path]


def _ptr(a):
    return a.ctypes.data_as(_f32p)


def oracle_log_mel(pcm: np.ndarray, filters: np.ndarray, n_threads: int = 1, return_raw: bool = False):
    """Restatement of whisper.cpp:109-216 for one clip. Returns [80][n/160] f32 (and the raw log10)."""
    build()
    lib = _lib(ORACLE_SO)
    pcm = np.ascontiguousarray(pcm, dtype=np.float32)
    filters = np.ascontiguousarray(filters, dtype=np.float32)
    n_mel = filters.shape[0]
    n_len = pcm.shape[0] // 160
    out = np.empty((n_mel, n_len), dtype=np.float32)
    raw = np.empty((n_mel, n_len), dtype=np.float32)
    rc = lib.oracle_log_mel(_ptr(pcm), ctypes.c_int(pcm.shape[0]), 400, 160, n_mel, int(n_threads),
                            _ptr(filters), _ptr(out), _ptr(raw))
    assert rc == 0
    return (out, raw) if return_raw else out


def oracle_fft(x: np.ndarray) -> np.ndarray:
    build()
    lib = _lib(ORACLE_SO)
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(2 * x.shape[0], dtype=np.float32)
    lib.oracle_fft(_ptr(x), ctypes.c_int(x.shape[0]), _ptr(out))
    return out


def oracle_argmax_last(x: np.ndarray) -> int:
    build()
    lib = _lib(ORACLE_SO)
    lib.oracle_argmax_last.restype = ctypes.c_longlong
    x = np.ascontiguousarray(x, dtype=np.float32)
    return int(lib.oracle_argmax_last(_ptr(x), ctypes.c_longlong(x.shape[0]), None))


def ref_log_mel(pcm: np.ndarray, filters: np.ndarray, n_threads: int = 0) -> np.ndarray:
    """The reference's own log_mel_spectrogram (n_threads=0 -> hardware_concurrency, as EncDec does)."""
    lib = _lib(REF_SO)
    pcm = np.ascontiguousarray(pcm, dtype=np.float32)
    filters = np.ascontiguousarray(filters, dtype=np.float32)
    out = np.empty((80, pcm.shape[0] // 160), dtype=np.float32)
    rc = lib.ref_log_mel(_ptr(pcm), ctypes.c_int(pcm.shape[0]), int(n_threads), _ptr(filters), _ptr(out))
    assert rc == 0
    return out


def ref_fft(x: np.ndarray) -> np.ndarray:
    lib = _lib(REF_SO)
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(2 * x.shape[0], dtype=np.float32)
    lib.ref_fft(_ptr(x), ctypes.c_int(x.shape[0]), _ptr(out))
    return out


def ref_hardware_concurrency() -> int:
    return int(_lib(REF_SO).ref_hardware_concurrency())


class RefVocab:
    """The reference's Reader / decode / language_id on a blob (whisper.cpp:519-665)."""

    def __init__(self, blob: bytes, multilingual: bool):
        self.lib = _lib(REF_SO)
        self._blob = ctypes.create_string_buffer(blob, len(blob))
        filt = np.empty(80 * 201, dtype=np.float32)
        n_mel, n_fft = ctypes.c_int(), ctypes.c_int()
        sp = (ctypes.c_int * 8)()
        head = ctypes.cast(ctypes.byref(self._blob, 8), ctypes.c_char_p)  # skip u64 (whisper.cpp:676-681)
        self.n_vocab = self.lib.ref_read_blob(head, int(multilingual), _ptr(filt), ctypes.byref(n_mel),
                                              ctypes.byref(n_fft), sp)
        self.filters = filt.reshape(n_mel.value, n_fft.value)
        keys = ["eot", "sot", "translate", "transcribe", "prev", "solm", "not", "beg"]
        self.special = dict(zip(keys, list(sp)))

    def token(self, i: int):
        buf = ctypes.create_string_buffer(512)
        n = self.lib.ref_token(int(i), buf, 512)
        return None if n < 0 else buf.raw[:n]

    def decode(self, ids, omit_special: bool = False) -> bytes:
        arr = (ctypes.c_int64 * len(ids))(*[int(i) for i in ids])
        buf = ctypes.create_string_buffer(1 << 16)
        n = self.lib.ref_decode(arr, len(ids), int(omit_special), buf, 1 << 16)
        assert n >= 0
        return buf.raw[:n]


def ref_language_id(code: str) -> int:
    return int(_lib(REF_SO).ref_language_id(code.encode()))


def ref_remove_extra_spaces(s: str) -> str:
    buf = ctypes.create_string_buffer(len(s.encode()) + 8)
    n = _lib(REF_SO).ref_remove_extra_spaces(s.encode(), buf, len(s.encode()) + 8)
    return buf.raw[:n].decode()


def ref_wav_read_legacy(path: str, cap: int = 1 << 24) -> np.ndarray:
    out = np.zeros(cap, dtype=np.float32)
    n = _lib(REF_SO).ref_wav_read_legacy(path.encode(), _ptr(out), cap)
    return out[:min(n, cap)].copy()


def ref_wav_read(path: str, cap: int = 1 << 24) -> np.ndarray:
    out = np.zeros(cap, dtype=np.float32)
    n = _lib(REF_SO).ref_wav_read(path.encode(), _ptr(out), cap)
    return out[:min(n, cap)].copy()


def ref_lang_code(i: int) -> str:
    lib = _lib(REF_SO)
    lib.ref_lang_code.restype = ctypes.c_char_p
    return lib.ref_lang_code(int(i)).decode()


def ref_language_count() -> int:
    return int(_lib(REF_SO).ref_language_count())
