"""GPU probe for the persistent decoder: small cases first, each in its own subprocess with a timeout."""
import os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def case(size, n):
    import numpy as np
    from tests.gpu_util import weight_file
    from whisper_tflite_b200 import capi, synth
    path, dims, W = weight_file(size)
    pcm = synth.synth_pcm(n, seed=21)
    os.environ["WT_DECODER"] = "legacy"
    old = capi.Engine(path, max_batch=n, keep_logits=n <= 8)
    del os.environ["WT_DECODER"]
    new = capi.Engine(path, max_batch=n, keep_logits=n <= 8)
    io, lo = old.transcribe_tokens(pcm)
    t = time.time(); i2, l2 = new.transcribe_tokens(pcm); dt = time.time() - t
    print(size, n, "launches old/new", old.launch_count, new.launch_count, "time new %.3f s" % dt)
    print(" ids old", io[0][:12].tolist(), "\n ids new", i2[0][:12].tolist(), "lens", l2[:4].tolist())
    print(" first-token agreement", float(np.mean(io[:, 4] == i2[:, 4])), "full agreement", float(np.mean(io == i2)))
    if n <= 8:
        for s in (0, 1, 26):
            a, b = old.step_logits(s, n), new.step_logits(s, n)
            print("  step", s, "max|dlogit|/max", float(np.abs(a - b).max() / np.abs(a).max()))

if __name__ == "__main__":
    if len(sys.argv) > 1:
        case(sys.argv[1], int(sys.argv[2]))
    else:
        for size, n in (("micro", 1), ("micro", 5), ("tiny", 3), ("tiny", 64), ("base", 16)):
            r = subprocess.run([sys.executable, __file__, size, str(n)], timeout=240)
            print("rc", r.returncode, flush=True)
            if r.returncode != 0:
                break
