"""Decoder launch time alone (tiny, B clips, given clips per cluster), stage profiler.
usage: [WT_DEC_CLUSTER=8] [WT_DEC_KV_AHEAD=n] python tools/dec_time.py [size] [B] [clips per cluster]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from whisper_tflite_b200 import capi, synth  # noqa: E402

size = sys.argv[1] if len(sys.argv) > 1 else "tiny"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
bench.MODEL = size
eng = capi.Engine(bench.ensure_weights(size), max_batch=B)
if len(sys.argv) > 3:
    eng.set_decoder_group(int(sys.argv[3]))
pcm = torch.from_numpy(synth.synth_pcm(B, seed=0)).cuda()
ids = torch.empty(B, 31, dtype=torch.int32, device="cuda")
lens = torch.empty(B, dtype=torch.int32, device="cuda")
for _ in range(2):
    eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
eng.profile(True)
for _ in range(3):
    eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
st = eng.profile_read()
print(f"cluster={os.environ.get('WT_DEC_CLUSTER', '16')} kv_ahead={os.environ.get('WT_DEC_KV_AHEAD', 'default')} group={sys.argv[3] if len(sys.argv) > 3 else 'auto'}: "
      f"decoder {st['dec_persistent'][0] / 3:.3f} ms; ids checksum {int(ids.sum().item())}")
