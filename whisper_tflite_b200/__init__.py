"""whisper.tflite_b200 — B200-native Whisper inference path behind the reference's whisper.h surface.

Only what the hot path needs lives here: ``csrc/`` (CUDA kernels + the C ABI + the host-side C++ mirror
of whisper.h), ``capi`` (ctypes binding of include/whisper_b200.h), ``synth`` (synthetic assets) and
``convert`` (HF <-> engine weight names).
"""
from . import synth  # noqa: F401

__all__ = ["synth", "capi", "convert", "build"]
