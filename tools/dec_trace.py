"""Per-phase time of the persistent decoder (CTA 0's view, %globaltimer) for the bench workload.
usage: python tools/dec_trace.py [size] [B] [clips per cluster] [fine]
`fine` builds and loads the -DDS_TRACE_FINE=1 variant of the library (LayerNorm / MMA rounds / epilogue pushes / fence /
barrier split inside the projection phases)."""
import os
import sys

os.environ["WT_DEC_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 4 and sys.argv[4] == "fine":
    from whisper_tflite_b200 import build as _b
    os.environ["WT_LIB"] = _b.build_cuda_variant("trace", ["-DDS_TRACE_FINE=1"])
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
