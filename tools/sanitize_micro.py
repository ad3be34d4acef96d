"""One micro-model transcribe (2 clips, bf16 production kernels) for compute-sanitizer:
    compute-sanitizer --tool racecheck|synccheck|initcheck python tools/sanitize_micro.py [size] [n_clips]
(ONE tool per gpurun call: B200_PROFILING.md.)"""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from whisper_tflite_b200 import capi, synth  # noqa: E402

size = sys.argv[1] if len(sys.argv) > 1 else "micro"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dims = synth.DIMS[size]
path = os.path.join(tempfile.gettempdir(), f"sanitize_{size}.b200w")
if not os.path.exists(path):
    synth.write_weight_file(path, dims, synth.init_weights(dims, seed=0))
eng = capi.Engine(path, max_batch=n)
ids, lens = eng.transcribe_tokens(synth.synth_pcm(n, seed=0))
print("lens", lens.tolist(), "launches", eng.launch_count)
eng.close()
