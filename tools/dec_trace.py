"""Per-phase time of the persistent decoder (CTA 0's view, %globaltimer) for the bench workload.
usage: WT_DEC_TRACE=1 python tools/dec_trace.py [size] [B]"""
import os
import sys

os.environ["WT_DEC_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from whisper_tflite_b200 import capi, synth  # noqa: E402

size = sys.argv[1] if len(sys.argv) > 1 else "tiny"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
bench.MODEL = size
eng = capi.Engine(bench.ensure_weights(size), max_batch=B)
pcm = torch.from_numpy(synth.synth_pcm(B, seed=0)).cuda()
ids = torch.empty(B, 31, dtype=torch.int32, device="cuda")
lens = torch.empty(B, dtype=torch.int32, device="cuda")
for _ in range(2):
    eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
eng.decoder_trace(reset=True)
eng.profile(True)
eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
st = eng.profile_read()
eng.profile(False)
print("stage ms:", {k: round(v[0], 3) for k, v in st.items() if v[1]})
tr = eng.decoder_trace()
tot = sum(v[0] for k, v in tr.items() if ":" not in k)
print(f"decoder trace {size} B={B}: total {tot/1e3:.3f} ms")
for k, (t, w) in tr.items():
    print(f"  {k:16s} {t:10.1f} us  ({100*t/tot:5.1f}%)  barrier wait {w:9.1f} us")
