"""Untraced time of the decoder stage (CUDA events around the launch) for the bench workload: min/median of N passes.
usage: python tools/dec_time.py [size] [B] [passes]"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from whisper_tflite_b200 import capi, synth  # noqa: E402

size = sys.argv[1] if len(sys.argv) > 1 else "tiny"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
N = int(sys.argv[3]) if len(sys.argv) > 3 else 7
bench.MODEL = size
eng = capi.Engine(bench.ensure_weights(size), max_batch=B)
pcm = torch.from_numpy(synth.synth_pcm(B, seed=0)).cuda()
ids = torch.empty(B, 31, dtype=torch.int32, device="cuda")
lens = torch.empty(B, dtype=torch.int32, device="cuda")
for _ in range(3):
    eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
ts = []
for _ in range(N):
    eng.profile(True)
    eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
    st = eng.profile_read()
    eng.profile(False)
    ts.append(st["dec_persistent"][0])
ts.sort()
clk = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm", "--format=csv,noheader"], capture_output=True,
                     text=True).stdout.strip()
print(f"decoder {size} B={B}: min {ts[0]:.3f} ms  median {ts[len(ts)//2]:.3f} ms  (sm clocks now/max: {clk})  ids sum {int(ids.sum())}")
