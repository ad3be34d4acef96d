"""Runs N transcribe passes of the bench workload (whisper-tiny, 64 clips, bf16) with nothing else —
the command ncu wraps for launch lists and full captures (B200_PROFILING.md recipe)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from whisper_tflite_b200 import capi, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
size = sys.argv[2] if len(sys.argv) > 2 else "tiny"
B = int(sys.argv[3]) if len(sys.argv) > 3 else 64
bench.MODEL = size
# the bench's arrangement: decoder clusters of 8 CTAs x 16 clips (ONE_STEP_CLUSTER / ONE_STEP_GROUP override)
eng = capi.Engine(bench.ensure_weights(size), max_batch=B, decoder_cluster=int(os.environ.get("ONE_STEP_CLUSTER", "8")))
eng.set_decoder_group(int(os.environ.get("ONE_STEP_GROUP", "16")))
init_launches = eng.launch_count
pcm = torch.from_numpy(synth.synth_pcm(B, seed=0)).cuda()
ids = torch.empty(B, 31, dtype=torch.int32, device="cuda")
lens = torch.empty(B, dtype=torch.int32, device="cuda")
for _ in range(n):
    eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
eng.synchronize()
print("lens", lens[:4].tolist(), "launches", eng.launch_count - init_launches, "init", init_launches)
