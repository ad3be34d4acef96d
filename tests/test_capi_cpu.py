"""CPU-side checks of the boundary: the C-ABI library loads, exports every symbol include/whisper_b200.h
declares, and refuses to run without a GPU (no compute calls here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from whisper_tflite_b200 import build, capi
    build.build_cuda()
    return capi.load_library()


def test_header_symbols_exported(lib):
    from whisper_tflite_b200 import capi
    hdr = open(os.path.join(ROOT, "include", "whisper_b200.h")).read()
    declared = set(re.findall(r"WT_API\s+[\w\s\*]+?\b(wt_\w+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(capi.EXPORTED)
    for s in declared:
        assert getattr(lib, s) is not None


def test_no_cpu_fallback(lib):
    import torch
    from whisper_tflite_b200 import capi
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(capi.WhisperB200Error, match="no CUDA device|CUDA"):
        capi.Engine(None)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "whisper_tflite_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f
                assert "oracle/" not in src or f == "synth.py", f


def test_header_is_plain_c():
    """include/whisper_b200.h is the drop-in boundary: it must compile as C99 (no C++ or torch types in the signatures)."""
    import shutil
    import subprocess
    cc = shutil.which("gcc") or shutil.which("cc")
    if not cc:
        pytest.skip("no C compiler")
    hdr = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "whisper_b200.h")
    r = subprocess.run([cc, "-fsyntax-only", "-x", "c", "-std=c99", "-Wall", "-Wextra", "-Werror", hdr], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
