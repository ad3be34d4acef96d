"""Two engines on two streams, batches alternating between them: does step i+1's mel/encoder overlap step i's decoder?
usage: python tools/pipe_probe.py [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from whisper_tflite_b200 import capi, synth  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
B = 64
w = bench.ensure_weights("tiny")
for n_eng in (1, 2, 3):
    engs = [capi.Engine(w, max_batch=B) for _ in range(n_eng)]
    streams = [torch.cuda.Stream() for _ in range(n_eng)]
    for e, s in zip(engs, streams):
        e.set_stream(s.cuda_stream)
    pcm = [torch.from_numpy(synth.synth_pcm(B, seed=i)).cuda() for i in range(2)]
    ids = [torch.empty(B, 31, dtype=torch.int32, device="cuda") for _ in range(n_eng)]
    lens = [torch.empty(B, dtype=torch.int32, device="cuda") for _ in range(n_eng)]

    def run(k):
        for i in range(k):
            j = i % n_eng
            engs[j].transcribe_tokens(pcm[i & 1], ids_out=ids[j], lens_out=lens[j])

    run(2 * n_eng)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(streams[0])
    for s in streams[1:]:
        s.wait_event(a)
    run(steps)
    for s in streams[1:]:
        streams[0].wait_stream(s)
    b.record(streams[0])
    torch.cuda.synchronize()
    print(f"{n_eng} engine(s): {a.elapsed_time(b) / steps:.3f} ms per step of {B} clips; lens ok {bool((lens[0] >= 5).all())}")
    for e in engs:
        e.close()
