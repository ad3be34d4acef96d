"""TEST INFRASTRUCTURE — builds the CPU-emulated kernel libraries (g++ only, no CUDA needed)."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(os.path.dirname(os.path.dirname(HERE)), "whisper_tflite_b200", "csrc")


def build(name: str) -> str:
    """name in {"mel", "simt"} -> path of libemu_<name>.so (rebuilt when any source is newer)."""
    so = os.path.join(HERE, f"libemu_{name}.so")
    srcs = [os.path.join(HERE, "cuda_emu.cpp"), os.path.join(HERE, f"emu_{name}.cpp")]
    deps = srcs + [os.path.join(HERE, "cuda_emu.h")] + [os.path.join(CSRC, f) for f in os.listdir(CSRC)
                                                         if f.endswith((".cuh", ".h"))]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so] + srcs + ["-lpthread"])
    return so
