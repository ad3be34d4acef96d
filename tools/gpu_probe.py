"""Development probe run on the GPU box: exercises each layer of the CUDA path in its own subprocess
(with a timeout, so a deadlocked kernel cannot hang the box) and prints parity numbers.

    python tools/gpu_probe.py [stage ...]      stages: mel gemm f32micro bf16micro f32tiny bf16tiny
"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def stage_mel():
    import numpy as np
    from oracle import mel as omel
    from whisper_tflite_b200 import capi, synth
    from tests.golden.make_mel_golden import golden_inputs
    omel.build()
    F = synth.mel_filters()
    eng = capi.Engine(None, max_batch=8)
    pcm = golden_inputs()
    out = eng.log_mel(pcm)
    for c in range(pcm.shape[0]):
        ref = omel.oracle_log_mel(pcm[c], F, 8)
        print(f"mel clip{c}: max|d|={np.abs(out[c]-ref).max():.3e} range={ref.max()-ref.min():.3f}")
    short = pcm[:2, :160 * 70 + 13].copy()
    o2 = eng.log_mel(short)
    ref = omel.oracle_log_mel(short[0], F, 4)
    print("ragged:", o2.shape, np.abs(o2[0] - ref).max())


def stage_gemm():
    import torch
    from whisper_tflite_b200 import capi
    torch.manual_seed(0)
    dev = "cuda"
    for (M, N, K, flags) in [(128, 64, 64, 0), (128, 128, 128, 0), (64, 384, 384, 0), (300, 1152, 384, 0),
                             (3000, 384, 1536, 0), (1500, 1536, 384, 2), (64, 51864, 384, 0), (257, 384, 240, 0),
                             (3000, 768, 384, 0), (4096, 512, 512, 0)]:
        a = (torch.randn(M, K, device=dev) * 0.5).bfloat16()
        w = (torch.randn(N, K, device=dev) * 0.1).bfloat16()
        b = torch.randn(N, device=dev)
        c = torch.zeros(M, N, device=dev, dtype=torch.bfloat16)
        capi.debug_gemm_bf16(a, w, b, c, flags)
        torch.cuda.synchronize()
        ref = a.float() @ w.float().T + b
        if flags & 2:
            ref = torch.nn.functional.gelu(ref)
        err = (c.float() - ref).abs().max().item()
        print(f"gemm M={M} N={N} K={K} flags={flags}: max|d|={err:.4f} ref max={ref.abs().max().item():.2f} "
              f"{'OK' if err < 0.02 * ref.abs().max().item() + 0.02 else 'FAIL'}")


def _model(size, precision, n_clips=2):
    import numpy as np
    from oracle import mel as omel, model_ref
    from whisper_tflite_b200 import capi, synth
    omel.build()
    dims = synth.DIMS[size]
    W = synth.init_weights(dims, seed=0)
    path = f"/tmp/wt_{size}.b200w"
    if not os.path.exists(path):
        synth.write_weight_file(path, dims, W)
    F = synth.mel_filters()
    pcm = synth.synth_pcm(n_clips, seed=0)
    eng = capi.Engine(path, max_batch=n_clips, precision=precision, keep_logits=True)
    mel = eng.log_mel(pcm)
    states = eng.encode(None, n_clips=n_clips)
    ids, lens = eng.decode_greedy(n_clips)
    for c in range(n_clips):
        mref = omel.oracle_log_mel(pcm[c], F, 8)
        xa = model_ref.encoder_forward(W, dims, mref)
        e = np.abs(states[c] - xa).max() / np.abs(xa).max()
        ids_ref, rows = model_ref.greedy_decode(W, dims, xa, list(capi.REFERENCE_PROMPT), capi.EOT_MULTILINGUAL,
                                                return_logits=True)
        n = int(lens[c])
        lg0 = eng.step_logits(0, n_clips)[c]
        le = np.abs(lg0 - rows[0]).max() / np.abs(rows[0]).max()
        same = list(ids[c][:n]) == ids_ref
        print(f"{size} prec={precision} clip{c}: enc rel-to-max err {e:.3e}; step0 logits rel err {le:.3e}; "
              f"len {n} vs {len(ids_ref)}; tokens identical: {same}")
        if not same:
            print("   got", list(ids[c][:n]))
            print("   ref", ids_ref)
    ids2, lens2 = eng.transcribe_tokens(pcm)
    print("transcribe_tokens == staged:", bool((ids2 == ids).all() and (lens2 == lens).all()), "launches", eng.launch_count)


STAGES = {
    "mel": stage_mel,
    "gemm": stage_gemm,
    "f32micro": lambda: _model("micro", 1),
    "bf16micro": lambda: _model("micro", 0),
    "f32tiny": lambda: _model("tiny", 1),
    "bf16tiny": lambda: _model("tiny", 0),
}

if __name__ == "__main__":
    if len(sys.argv) == 3 and sys.argv[1] == "--run":
        STAGES[sys.argv[2]]()
        sys.exit(0)
    stages = sys.argv[1:] or list(STAGES)
    for s in stages:
        t = time.time()
        print(f"===== {s} =====", flush=True)
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--run", s], timeout=420, cwd=ROOT)
            print(f"===== {s}: exit {r.returncode} in {time.time()-t:.1f}s", flush=True)
        except subprocess.TimeoutExpired:
            print(f"===== {s}: TIMEOUT (killed)", flush=True)
