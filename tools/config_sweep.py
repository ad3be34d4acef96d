"""BASELINE.json configs[3..4] shapes (random-init base / small) next to the headline tiny config: one engine, steps back
to back, device-resident PCM. Prints one JSON line per (model, batch). usage: python tools/config_sweep.py [out.jsonl]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from whisper_tflite_b200 import capi, synth  # noqa: E402

out = open(sys.argv[1], "w") if len(sys.argv) > 1 else None
for size, B in (("tiny", 64), ("tiny", 256), ("base", 256), ("small", 64), ("small", 256)):
    bench.MODEL = size
    eng = capi.Engine(bench.ensure_weights(size), max_batch=B)
    stream = torch.cuda.Stream()
    eng.set_stream(stream.cuda_stream)
    pcm = torch.from_numpy(synth.synth_pcm(B, seed=0)).cuda()
    ids = torch.empty(B, 31, dtype=torch.int32, device="cuda")
    lens = torch.empty(B, dtype=torch.int32, device="cuda")
    for _ in range(3):
        eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = 5
    a.record(stream)
    for _ in range(steps):
        eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
    b.record(stream)
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    eng.profile(True)
    eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
    st = eng.profile_read()
    eng.profile(False)
    line = {"model": size, "clips": B, "ms_per_step": round(ms, 3), "rtfx": round(B * 30.0 / (ms / 1e3), 1),
            "stages_ms": {k: round(v[0], 3) for k, v in st.items() if v[1]}, "lens_ok": bool((lens >= 5).all())}
    print(json.dumps(line), flush=True)
    if out:
        out.write(json.dumps(line) + "\n")
    eng.close()
    del pcm
    torch.cuda.empty_cache()
