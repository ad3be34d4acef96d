"""The streaming persistent decoder (csrc/decoder_stream_sm100.cuh) and the serving pool around it, through the C ABI.

Parity against the oracle and against the per-kernel decoder is in test_gpu_model.py (every bf16 test there runs this
kernel); here: invariance under the clips-per-cluster setting (wt_set_decoder_group), the early-EOT exit (the producer
warp has copies in flight that nobody will consume), ragged last groups, and capi.EnginePool against a single engine.
"""
import numpy as np
import pytest

from tests.gpu_util import weight_file
from whisper_tflite_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("size,n_clips", [("micro", 19), ("tiny", 21)])
def test_decoder_group_size_does_not_change_tokens(size, n_clips):
    """Which CTA owns which output columns / (clip, head) pairs changes with the group size; the arithmetic per element
    (K order, chunk order, merge order) does not — token ids are identical."""
    from whisper_tflite_b200 import capi
    path, dims, W = weight_file(size)
    pcm = synth.synth_pcm(n_clips, seed=5)
    eng = capi.Engine(path, max_batch=n_clips)
    ref_ids, ref_lens = eng.transcribe_tokens(pcm)
    assert np.all(ref_lens >= 5)
    for group in (16, 9, 8):   # 2 / 3 clusters with a ragged last group
        eng.set_decoder_group(group)
        ids, lens = eng.transcribe_tokens(pcm)
        np.testing.assert_array_equal(ids, ref_ids, err_msg=f"group {group}")
        np.testing.assert_array_equal(lens, ref_lens)
    with pytest.raises(capi.WhisperB200Error):
        eng.set_decoder_group(17)
    eng.close()


def test_early_eot_exits_and_engine_stays_usable():
    """All clips hit EOT after 3 generated tokens: every cluster leaves the loop with weight / K-V copies still in flight
    (the producer drains them). The engine must return padded ids and keep working."""
    from whisper_tflite_b200 import capi
    path, dims, W = weight_file("micro")
    pcm = synth.synth_pcm(1, seed=11)
    eng = capi.Engine(path, max_batch=1)
    full, full_lens = eng.transcribe_tokens(pcm)
    eot = int(full[0][4 + 2])
    first = int(np.argmax(full[0][4:] == eot))          # first generated position holding that token
    ids, lens = eng.transcribe_tokens(pcm, eot=eot)
    assert lens[0] == 4 + first + 1
    np.testing.assert_array_equal(ids[0][:lens[0]], full[0][:lens[0]])
    assert np.all(ids[0][lens[0]:] == -1)
    again, again_lens = eng.transcribe_tokens(pcm)       # the next launch starts from clean barriers and rings
    np.testing.assert_array_equal(again, full)
    np.testing.assert_array_equal(again_lens, full_lens)
    eng.close()


def test_engine_pool_matches_single_engine():
    """capi.EnginePool: batches alternate between two engines on separate streams (one host thread each); results are
    those of one engine taking the batches one after the other."""
    from whisper_tflite_b200 import capi
    path, dims, W = weight_file("micro")
    batches = [synth.synth_pcm(6, seed=30 + i) for i in range(5)]
    one = capi.Engine(path, max_batch=6)
    want = [tuple(a.copy() for a in one.transcribe_tokens(b)) for b in batches]
    one.close()
    pool = capi.EnginePool(path, n_engines=2, max_batch=6)
    got = pool.transcribe_tokens_many(batches)
    for (ids, lens), (wi, wl) in zip(got, want):
        np.testing.assert_array_equal(ids, wi)
        np.testing.assert_array_equal(lens, wl)
    assert pool.launch_count > 0
    pool.close()


@pytest.mark.parametrize("size,n_clips", [("base", 3), ("small", 2)])
def test_base_and_small_dims(size, n_clips, oracle_built, filters):
    """BASELINE.json configs[3] / [4] architectures (random init): d = 512 (8 weight units per warp and stage, 8 heads)
    and d = 768 (2-stage ring next to the 99 KB hidden tile, 12 + 12 layers). Encoder states against the oracle, the
    streaming decoder against the per-kernel decoder on the same bf16 operands."""
    import os

    from oracle import model_ref
    from tests.gpu_util import rel_to_max
    from whisper_tflite_b200 import capi
    path, dims, W = weight_file(size)
    pcm = synth.synth_pcm(n_clips, seed=9)
    os.environ["WT_DECODER"] = "legacy"
    try:
        old = capi.Engine(path, max_batch=n_clips, keep_logits=True)
    finally:
        del os.environ["WT_DECODER"]
    new = capi.Engine(path, max_batch=n_clips, keep_logits=True)
    new.log_mel(pcm)
    states = new.encode(None, n_clips=n_clips)
    mel0 = oracle_built.oracle_log_mel(pcm[0], filters, 8)
    xa0 = model_ref.encoder_forward(W, dims, mel0)
    assert rel_to_max(states[0], xa0) <= 3e-2
    ids_new, lens_new = new.decode_greedy(n_clips)
    ids_old, lens_old = old.transcribe_tokens(pcm)
    assert np.all(lens_new >= 5)
    for c in range(n_clips):
        for s_ in range(27):
            a, b = old.step_logits(s_, n_clips)[c], new.step_logits(s_, n_clips)[c]
            err = np.abs(a - b).max()
            assert err <= 2e-2 * np.abs(a).max(), (c, s_, err)
            top2 = np.partition(a, -2)[-2:]
            if top2[1] - top2[0] > 4 * err:
                assert ids_old[c][4 + s_] == ids_new[c][4 + s_]
            else:
                break
    old.close()
    new.close()
