"""Shared helpers for the -m gpu parity tests (all calls go through the C ABI via capi)."""
import os
import tempfile

import numpy as np

from whisper_tflite_b200 import synth

_CACHE = {}


def weight_file(size: str, seed: int = 0):
    """-> (path, dims, weights dict). Files are cached in the temp dir for the session."""
    key = (size, seed)
    if key not in _CACHE:
        dims = synth.DIMS[size]
        W = synth.init_weights(dims, seed=seed)
        path = os.path.join(tempfile.gettempdir(), f"wt_test_{size}_{seed}.b200w")
        synth.write_weight_file(path, dims, W)
        _CACHE[key] = (path, dims, W)
    return _CACHE[key]


def rel_to_max(a, b):
    return float(np.abs(a - b).max() / max(1e-30, np.abs(b).max()))
