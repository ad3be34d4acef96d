"""Per-stage device times of one engine on the bench workload (whisper-tiny, 64 clips), engine-side CUDA events.

    python tools/stage_time.py [n_clips] [model]        env switches (WT_GEMM_LEAN, WT_DECODER ...) apply as usual
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from tests.gpu_util import weight_file
    from whisper_tflite_b200 import capi, synth
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    model = sys.argv[2] if len(sys.argv) > 2 else "tiny"
    path, dims, W = weight_file(model)
    eng = capi.Engine(path, max_batch=B)
    pcm = torch.from_numpy(synth.synth_pcm(B, seed=1)).cuda()
    ids = torch.empty(B, capi.MAX_TOKENS, dtype=torch.int32, device="cuda")
    lens = torch.empty(B, dtype=torch.int32, device="cuda")
    for _ in range(3):
        eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
    torch.cuda.synchronize()
    n = 5
    eng.profile(True)
    for _ in range(n):
        eng.transcribe_tokens(pcm, ids_out=ids, lens_out=lens)
    st = eng.profile_read()
    eng.profile(False)
    tot = 0.0
    for k, v in st.items():
        if v[1]:
            print(f"{k:18s} {v[0] / n:8.4f} ms  {v[1] / n:6.1f} launches")
            tot += v[0] / n
    print(f"{'total':18s} {tot:8.4f} ms")
    eng.close()


if __name__ == "__main__":
    main()
