"""Generates tests/golden/mel_golden.npz from the reference's OWN log_mel_spectrogram
(oracle/_ref/libwhisper_ref.so = /root/reference/whisper.tflite/whisper.cpp:109-216 compiled in
place). Run in the build container: `python tests/golden/make_mel_golden.py`.

Inputs are regenerated from seeds by `golden_inputs()` below (imported by the tests); stored are
every 37th frame of each clip's mel plus whole-array statistics, so the fixture stays small.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from whisper_tflite_b200 import synth  # noqa: E402

STRIDE = 37


def golden_inputs():
    """6 clips x 480000 f32: two BASELINE clips, silence, DC, a unit impulse train, full-scale noise."""
    n = synth.N_SAMPLES
    pcm = np.zeros((6, n), np.float32)
    pcm[0:2] = synth.synth_pcm(2, seed=0)
    pcm[3] = 0.25
    pcm[4, ::4001] = 1.0
    pcm[5] = np.random.RandomState(5).uniform(-1, 1, n).astype(np.float32)
    return pcm


def main():
    from oracle import mel as omel
    F = synth.mel_filters()
    pcm = golden_inputs()
    mels = np.stack([omel.ref_log_mel(p, F, 0) for p in pcm])
    out = os.path.join(ROOT, "tests", "golden", "mel_golden.npz")
    np.savez_compressed(out, sample=mels[:, :, ::STRIDE].copy(), stride=STRIDE,
                        mean=mels.mean(axis=(1, 2), dtype=np.float64), mx=mels.max(axis=(1, 2)),
                        mn=mels.min(axis=(1, 2)), abs_sum=np.abs(mels).sum(axis=(1, 2), dtype=np.float64))
    print("wrote", out, os.path.getsize(out), "bytes", mels.min(), mels.max())


if __name__ == "__main__":
    main()
