"""BASELINE.json configs[2]: one hour of 16 kHz mono audio (57.6 M samples) -> 120 windows of 30 s (shard.chunk_windows)
-> token ids, through the PCM-in / ids-out call (wt_transcribe_tokens, the Monolith path: whisper.cpp:692-738) with HOST
buffers, batches of <= 64 windows on an EnginePool; then the same stream repeated x32 (3 840 windows) so that the work
is long enough to read a rate from. Wall clock around the calls (they return synchronised). One JSON line per run.
usage: python tools/long_audio.py [out.jsonl] [repeat=32] [engines=3]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from whisper_tflite_b200 import capi, shard, synth  # noqa: E402

B = 64
out = open(sys.argv[1], "w") if len(sys.argv) > 1 and sys.argv[1] != "-" else None
repeat = int(sys.argv[2]) if len(sys.argv) > 2 else 32
hour = synth.synth_pcm(120, seed=7).reshape(-1)            # 57.6 M samples = 1 h
windows = shard.chunk_windows(hour)
n_win = windows.shape[0]
assert n_win == 120
# n_win + B rows so that every batch of the repeated stream is one contiguous slice
ring = torch.empty(n_win + B, shard.WINDOW, dtype=torch.float32).pin_memory()
ring[:n_win] = torch.from_numpy(windows)
ring[n_win:] = ring[:B]
n_eng = int(sys.argv[3]) if len(sys.argv) > 3 else 3
pool = capi.EnginePool(bench.ensure_weights("tiny"), n_engines=n_eng, max_batch=B)


def run(total):
    starts = list(range(0, total, B))
    ids = torch.full((len(starts), B, capi.MAX_TOKENS), -1, dtype=torch.int32).pin_memory()
    lens = torch.zeros(len(starts), B, dtype=torch.int32).pin_memory()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    futs = []
    for k, s in enumerate(starts):
        n = min(B, total - s)
        o = s % n_win
        futs.append(pool.submit(ring[o:o + n], ids_out=ids[k, :n], lens_out=lens[k, :n]))
    for f in futs:
        f.result()
    dt = time.perf_counter() - t0
    return dt, ids.reshape(-1, capi.MAX_TOKENS)[:total].numpy(), lens.reshape(-1)[:total].numpy()


run(n_win)                                                  # warm-up
dt1, ids1, lens1 = run(n_win)
line = {"workload": "1 h of audio = 120 windows, host PCM in, ids out, %d engines" % n_eng, "windows": n_win, "seconds": round(dt1, 5),
        "rtfx": round(n_win * 30.0 / dt1, 1), "lens_ok": bool((lens1 >= 5).all())}
print(json.dumps(line), flush=True)
if out:
    out.write(json.dumps(line) + "\n")
total = n_win * repeat
dt, ids, lens = run(total)
same = all(np.array_equal(ids[r * n_win:(r + 1) * n_win], ids1) and np.array_equal(lens[r * n_win:(r + 1) * n_win], lens1)
           for r in range(repeat))
rep = ids.reshape(repeat, n_win, -1)
n_diff = int((rep != ids1[None]).any(axis=2).sum())
# batches of 64 over a period of 120 windows: repeats r and r + 8 see the same batch composition and batch positions
period = 8
same_align = bool(repeat <= period or np.array_equal(rep[:repeat - period], rep[period:]))
line = {"workload": f"the same hour x{repeat}", "windows": total, "seconds": round(dt, 5), "rtfx": round(total * 30.0 / dt, 1),
        "ids_identical_to_single_pass": bool(same), "windows_differing_from_single_pass": n_diff,
        "identical_at_equal_batch_alignment": same_align}
print(json.dumps(line), flush=True)
if out:
    out.write(json.dumps(line) + "\n")
pool.close()
