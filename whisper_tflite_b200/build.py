"""In-tree build of the native pieces (explicit nvcc/g++ commands; no JIT cache, so the .so files travel
with the repo snapshot to the GPU box).

  csrc/libwhisper_b200.so      CUDA kernels + C ABI (include/whisper_b200.h), sm_100a only
  csrc/host/libwhisper_host.so host-side C++ mirror of the reference's whisper.h over the C ABI
  csrc/host/encdec, minimal    the reference's CLI drivers rebuilt on the mirror
"""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(CSRC, "host")
LIB = os.path.join(CSRC, "libwhisper_b200.so")
HOST_LIB = os.path.join(HOST, "libwhisper_host.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden", "-shared"]


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _nvcc() -> str:
    for c in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found: the CUDA library cannot be built (there is no CPU fallback)")


def build_cuda(force: bool = False, verbose: bool = False) -> str:
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h"))]
    deps.append(os.path.join(os.path.dirname(HERE), "include", "whisper_b200.h"))
    if force or _newer(LIB, deps):
        cmd = [_nvcc()] + NVCC_FLAGS + os.environ.get("WT_NVCC_EXTRA", "").split() + (["-Xptxas", "-v"] if verbose else []) + \
              ["-o", LIB, os.path.join(CSRC, "engine.cu"), "-lcudart"]
        subprocess.check_call(cmd)
    return LIB


def build_cuda_variant(suffix: str, extra_flags) -> str:
    """A second in-tree library from the same sources with extra nvcc flags (profiling variants: -DDS_TRACE_FINE=1,
    -DDS_PROBE); selected at run time with WT_LIB=<path>."""
    out = os.path.join(CSRC, f"libwhisper_b200_{suffix}.so")
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cuh", ".h"))]
    if _newer(out, deps):
        subprocess.check_call([_nvcc()] + NVCC_FLAGS + list(extra_flags) + ["-o", out, os.path.join(CSRC, "engine.cu"), "-lcudart"])
    return out


def build_host(force: bool = False) -> str:
    if not os.path.isdir(HOST):
        return ""
    srcs = [os.path.join(HOST, f) for f in sorted(os.listdir(HOST)) if f.endswith(".cpp") and not f.startswith(("app_", "jni_"))]
    deps = srcs + [os.path.join(HOST, f) for f in os.listdir(HOST) if f.endswith(".h")] + [LIB]
    if srcs and (force or _newer(HOST_LIB, deps)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", HOST_LIB] + srcs +
                              ["-I", os.path.join(os.path.dirname(HERE), "include"), "-L", CSRC, "-lwhisper_b200",
                               "-Wl,-rpath,$ORIGIN/..", "-lpthread"])
    # the JNI bridge has no JDK to link against here: compile-check it against the declaration-only stub
    jni = os.path.join(HOST, "jni_whisper.cpp")
    jni_o = os.path.join(HOST, "jni_whisper.o")
    if os.path.exists(jni) and (force or _newer(jni_o, [jni, os.path.join(HOST, "whisper.h")])):
        jdk = os.environ.get("JAVA_HOME")
        inc = (["-I", os.path.join(jdk, "include"), "-I", os.path.join(jdk, "include", "linux")] if jdk and
               os.path.exists(os.path.join(jdk, "include", "jni.h")) else ["-I", os.path.join(HOST, "jni_stub")])
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-c", jni, "-o", jni_o, "-I", HOST] + inc)
    for app in ("encdec", "minimal"):
        src = os.path.join(HOST, f"app_{app}.cpp")
        exe = os.path.join(HOST, app)
        if os.path.exists(src) and (force or _newer(exe, [src, HOST_LIB])):
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-o", exe, src, "-I", HOST,
                                   "-I", os.path.join(os.path.dirname(HERE), "include"), "-L", HOST, "-lwhisper_host",
                                   "-L", CSRC, "-lwhisper_b200", "-Wl,-rpath,$ORIGIN:$ORIGIN/..", "-lpthread"])
    return HOST_LIB


def build_all(force: bool = False) -> None:
    build_cuda(force)
    build_host(force)


if __name__ == "__main__":
    import sys
    build_all(force="--force" in sys.argv)
    print("built", LIB)
