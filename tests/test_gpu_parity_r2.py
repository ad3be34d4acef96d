"""Round-2 parity gates for the PRODUCTION bf16 kernels (gemm_bf16_sm100_kernel, attention3_sm100_kernel,
decoder_stream_kernel), through the C ABI:

  * bitwise reproducibility of one tiny/64 batch repeated >= 200 times on a three-engine pool (ids AND digests of the
    encoder states / cross K/V) — one window -> one id list (whisper.cpp:367-400);
  * the streaming decoder against the oracle at EVERY decode step (teacher-forced with the oracle's own prefix, so a
    low-margin flip at step s cannot hide steps s+1..26), tiny and base;
  * BASELINE's configuration (tiny, 64 clips): encoder states and step logits of sampled clips (first, last, one per
    decoder cluster) against the oracle, and clip c inside the batch of 64 == clip c alone;
  * the argmax rule of whisper.cpp:346-361 on the device, in all three argmax paths: exact ties -> LAST index,
    NaN never wins unless it is element 0.
"""
import os
import threading

import numpy as np
import pytest

from oracle import model_ref
from tests.gpu_util import rel_to_max, weight_file
from whisper_tflite_b200 import synth

pytestmark = pytest.mark.gpu
PROMPT = [50258, 50261, 50359, 50363]
EOT = 50257
BF16_TOL = 3e-2   # rel-to-max: bf16 operands, f32 accumulate, bf16 residual stream (encoder) through 4+4 layers


def _oracle_clip(size, pcm_clip, oracle_built, filters, W=None, dims=None):
    if W is None:
        _, dims, W = weight_file(size)
    mel = oracle_built.oracle_log_mel(pcm_clip, filters, 8)
    xa = model_ref.encoder_forward(W, dims, mel)
    ids, rows = model_ref.greedy_decode(W, dims, xa, PROMPT, EOT, return_logits=True)
    return xa, ids, rows


def test_bitwise_reproducible_under_engine_pool():
    """The arrangement bench.py times: three engines on one GPU, one host thread each, all busy. The SAME 64 clips go
    through every engine 70 times (210 passes): token ids and the digests of the encoder states and of a cross-K/V
    stream must be identical every time. (Round 1 saw 0.7 % of windows differ: a write-after-read race on the GEMM
    epilogue's bias staging — gemm_sm100.cuh, GemmSmem::kBiasBytes.)"""
    from whisper_tflite_b200 import capi
    path, dims, W = weight_file("tiny")
    B, n_eng, reps = 64, 3, 70
    pcm = synth.synth_pcm(B, seed=77)
    engs = [capi.Engine(path, max_batch=B) for _ in range(n_eng)]
    for e in engs:
        e.set_decoder_group(16)
    results = [[] for _ in range(n_eng)]
    errors = []

    def work(k):
        try:
            for _ in range(reps):
                ids, lens = engs[k].transcribe_tokens(pcm)
                results[k].append((ids.copy(), lens.copy(), engs[k].debug_hash("enc_out", B),
                                   engs[k].debug_hash("cross_kv3", B), engs[k].debug_hash("x", B)))
        except Exception as ex:  # surfaced below: an exception in a thread must fail the test
            errors.append(ex)

    threads = [threading.Thread(target=work, args=(k,)) for k in range(n_eng)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    ids0, lens0, h_enc0, h_kv0, h_x0 = results[0][0]
    assert np.all(lens0 >= 5)
    bad = []
    for k in range(n_eng):
        assert len(results[k]) == reps
        for r, (ids, lens, h_enc, h_kv, h_x) in enumerate(results[k]):
            if h_x != h_x0 or h_enc != h_enc0 or h_kv != h_kv0 or not np.array_equal(ids, ids0) or not np.array_equal(lens, lens0):
                bad.append((k, r, h_x == h_x0, h_enc == h_enc0, h_kv == h_kv0, int((ids != ids0).any(axis=1).sum())))
    for e in engs:
        e.close()
    assert not bad, f"{len(bad)} of {n_eng * reps} passes differ (engine, repeat, x==, enc==, kv==, clips with other ids): {bad[:8]}"


@pytest.mark.parametrize("size,n_clips", [("tiny", 2), ("base", 1)])
def test_stream_decoder_teacher_forced_logits_every_step(size, n_clips, oracle_built, filters):
    """decoder_stream_kernel against the oracle at every step s = 0..26: the prompt is the oracle's own prefix
    ids[:4+s], so the kernel's logits at position 3+s (KV-cache append, self-attention over a growing cache, position
    embedding, cross-attention, logits) are compared with the oracle's row s whatever happened at earlier steps."""
    from whisper_tflite_b200 import capi
    path, dims, W = weight_file(size)
    pcm = synth.synth_pcm(n_clips, seed=0)
    eng = capi.Engine(path, max_batch=n_clips, keep_logits=True)
    eng.log_mel(pcm)
    states = eng.encode(None, n_clips=n_clips)
    worst = 0.0
    for c in range(n_clips):
        xa, ids_ref, rows = _oracle_clip(size, pcm[c], oracle_built, filters)
        assert rel_to_max(states[c], xa) <= BF16_TOL
        n_steps = rows.shape[0]
        assert n_steps >= 20      # random-init weights do not emit EOT early
        flips = 0
        for s in range(n_steps):
            prefix = ids_ref[:4 + s]
            ids, lens = eng.decode_greedy(n_clips, prompt=prefix, max_total_tokens=5 + s)
            lg = eng.step_logits(0, n_clips)[c]
            err = float(np.abs(lg - rows[s]).max() / np.abs(rows[s]).max())
            worst = max(worst, err)
            assert err <= BF16_TOL, (size, c, s, err)
            assert list(ids[c][:4 + s]) == prefix and lens[c] == 5 + s
            # the token it appends is the oracle's wherever the oracle's decision margin exceeds the logit error
            top2 = np.partition(rows[s], -2)[-2:]
            if top2[1] - top2[0] > 2 * np.abs(lg - rows[s]).max():
                assert ids[c][4 + s] == ids_ref[4 + s], (size, c, s)
            else:
                flips += 1
        print(f"{size} clip {c}: {n_steps} steps, worst rel-to-max logit error {worst:.2e}, {flips} low-margin steps")
    eng.close()


def test_tiny64_against_oracle_and_batch_invariance(oracle_built, filters):
    """BASELINE.json configs[1] itself: 64 clips through one call (750 GEMM row tiles per launch, per-clip cross-K/V
    launches, 4 decoder clusters x 16 clips). Sampled clips (first, last, one in every decoder cluster) against the
    oracle: encoder states, and the logits of every step whose prefix still equals the oracle's; then the same clips
    ALONE (batch of 1) must give the ids they gave inside the batch."""
    from whisper_tflite_b200 import capi
    path, dims, W = weight_file("tiny")
    B = 64
    pcm = synth.synth_pcm(B, seed=1000)   # bench.py's first PCM set on rank 0
    eng = capi.Engine(path, max_batch=B, keep_logits=True)
    eng.set_decoder_group(16)
    eng.log_mel(pcm)
    states = eng.encode(None, n_clips=B)
    ids, lens = eng.decode_greedy(B)
    assert np.all(lens >= 5)
    sampled = [0, 21, 38, 63]
    for c in sampled:
        xa, ids_ref, rows = _oracle_clip("tiny", pcm[c], oracle_built, filters)
        assert rel_to_max(states[c], xa) <= BF16_TOL, c
        compared = 0
        for s in range(rows.shape[0]):
            if list(ids[c][:4 + s]) != ids_ref[:4 + s]:
                break
            lg = eng.step_logits(s, B)[c]
            assert np.abs(lg - rows[s]).max() <= BF16_TOL * np.abs(rows[s]).max(), (c, s)
            compared += 1
        assert compared >= 1
    # fused-argmax production engine (no logits in HBM): same ids; and each sampled clip alone == inside the batch
    prod = capi.Engine(path, max_batch=B)
    prod.set_decoder_group(16)
    ids_p, lens_p = prod.transcribe_tokens(pcm)
    np.testing.assert_array_equal(ids_p, ids)
    np.testing.assert_array_equal(lens_p, lens)
    for c in sampled:
        i1, l1 = prod.transcribe_tokens(np.ascontiguousarray(pcm[c:c + 1]))
        np.testing.assert_array_equal(i1[0], ids[c], err_msg=f"clip {c} alone vs in the batch of 64")
        assert l1[0] == lens[c]
    # a ragged batch (last decoder cluster holds 5 clips): same ids for its clips
    i37, l37 = prod.transcribe_tokens(np.ascontiguousarray(pcm[:37]))
    np.testing.assert_array_equal(i37, ids[:37])
    eng.close()
    prod.close()


# ---------------------------------------------------------------------------------------------------
# whisper.cpp:346-361 on the device: `>=` scan => LAST maximal index; NaN never wins unless it is element 0
# ---------------------------------------------------------------------------------------------------
def _engines_for_all_argmax_paths(path, n_clips):
    """-> [(name, engine)]: the three argmax implementations the library has.
    stream: decoder_stream_kernel's ds_logits (production); epi_argmax: the tcgen05 GEMM's EPI_ARGMAX epilogue +
    argmax_partials_kernel (per-kernel bf16 decoder without kept logits); append: argmax_append_kernel over f32 logits
    (fp32 verification mode)."""
    from whisper_tflite_b200 import capi
    out = [("stream", capi.Engine(path, max_batch=n_clips))]
    os.environ["WT_DECODER"] = "legacy"
    try:
        out.append(("epi_argmax", capi.Engine(path, max_batch=n_clips)))
    finally:
        del os.environ["WT_DECODER"]
    out.append(("append", capi.Engine(path, max_batch=n_clips, precision=capi.PRECISION_FP32_VERIFY)))
    return out


def _write_weights(tag, dims, W):
    import tempfile
    path = os.path.join(tempfile.gettempdir(), f"wt_test_{tag}.b200w")
    synth.write_weight_file(path, dims, W)
    return path


def test_exact_ties_take_the_last_index_on_device(oracle_built, filters):
    """Two more vocabulary rows are made bit-identical to the row of the token that wins step 0 (tied embedding: their
    logits are identical in every path, the MMA accumulation order does not depend on the column). The largest of the
    three indices must come out — across lanes, warps, CTAs of a cluster (stream), N tiles (EPI_ARGMAX) and threads
    (argmax_append) — and the following steps must follow the oracle's `>=` rule too."""
    from whisper_tflite_b200 import capi
    path0, dims, W0 = weight_file("micro")
    n_clips = 2
    pcm = synth.synth_pcm(n_clips, seed=4)
    base = capi.Engine(path0, max_batch=n_clips)
    ids0, _ = base.transcribe_tokens(pcm)
    base.close()
    V = dims.n_vocab
    t0 = int(ids0[0][4])
    a, b = (t0 + V // 3) % V, (t0 + 2 * V // 3) % V
    assert len({t0, a, b} - set(PROMPT) - {EOT}) == 3
    W = {k: v.copy() for k, v in W0.items()}
    emb = W["decoder.token_embedding.weight"]
    emb[a] = emb[t0]
    emb[b] = emb[t0]
    want_first = max(t0, a, b)
    path = _write_weights("micro_ties", dims, W)
    ref_ids = []
    for c in range(n_clips):
        mel = oracle_built.oracle_log_mel(pcm[c], filters, 8)
        xa = model_ref.encoder_forward(W, dims, mel)
        ref_ids.append(model_ref.greedy_decode(W, dims, xa, PROMPT, EOT))
    assert ref_ids[0][4] == want_first     # the oracle restates the `>=` scan
    for name, eng in _engines_for_all_argmax_paths(path, n_clips):
        ids, lens = eng.transcribe_tokens(pcm)
        assert int(ids[0][4]) == want_first, (name, int(ids[0][4]), (t0, a, b))
        if name == "append":   # f32 path: every token is the oracle's
            for c in range(n_clips):
                assert list(ids[c][:lens[c]]) == ref_ids[c], name
        else:                  # bf16 paths: every later member of the tied triple must again be its largest index
            for c in range(n_clips):
                for tok in ids[c][4:lens[c]]:
                    assert int(tok) not in (min(t0, a, b), sorted((t0, a, b))[1]), (name, c, int(tok))
        eng.close()


@pytest.mark.parametrize("nan_row", [0, 12345])
def test_nan_logit_never_wins_unless_first(nan_row, oracle_built, filters):
    """A NaN embedding row makes that token's logit NaN at every step, in every path. whisper.cpp:348-357: NaN at
    index > 0 never satisfies `>=` (ids are those of the clean model); NaN at index 0 is the initial max_value and
    survives every comparison (token 0 is appended, its NaN embedding is fed back, and token 0 wins from then on)."""
    path0, dims, W0 = weight_file("micro")
    n_clips = 2
    pcm = synth.synth_pcm(n_clips, seed=6)
    W = {k: v.copy() for k, v in W0.items()}
    W["decoder.token_embedding.weight"][nan_row] = np.float32("nan")
    path = _write_weights(f"micro_nan{nan_row}", dims, W)
    clean = dict(_engines_for_all_argmax_paths(path0, n_clips))
    for name, eng in _engines_for_all_argmax_paths(path, n_clips):
        ids, lens = eng.transcribe_tokens(pcm)
        if nan_row == 0:
            assert np.all(lens == 31), name
            assert np.all(ids[:, 4:] == 0), (name, ids[:, 4:8])
        else:
            ids_c, lens_c = clean[name].transcribe_tokens(pcm)
            np.testing.assert_array_equal(ids, ids_c, err_msg=name)
            assert not np.any(ids[:, 4:] == nan_row)
        eng.close()
    for e in clean.values():
        e.close()
    # the oracle restates the same rule
    row = np.array([1.0, np.nan, 3.0, 3.0, np.nan], np.float32)
    assert model_ref.argmax_last(row) == 3
    row[0] = np.nan
    assert model_ref.argmax_last(row) == 0


def test_decoder_cluster_sizes_give_identical_tokens():
    """The pool's decoder arrangement (wt_config.decoder_cluster = 4 or 8 CTAs per 16-clip group, bench.py's default for
    the serving pool) against the single-call default (16-CTA clusters, the batch spread over every co-resident cluster):
    which CTA owns a column slice or a (clip, head) pair changes, the arithmetic per element does not — ids, lens and the
    step-0 logits are bit-identical, with a ragged last group (40 clips = 16 + 16 + 8)."""
    from whisper_tflite_b200 import capi
    path, dims, W = weight_file("tiny")
    n = 40
    pcm = synth.synth_pcm(n, seed=21)
    ref = None
    for cluster, group in ((0, 0), (16, 16), (8, 16), (4, 16), (4, 8)):
        eng = capi.Engine(path, max_batch=n, decoder_cluster=cluster, keep_logits=True)
        eng.set_decoder_group(group)
        ids, lens = eng.transcribe_tokens(pcm)
        lg = eng.step_logits(0, n).copy()
        eng.close()
        if ref is None:
            ref = (ids.copy(), lens.copy(), lg)
            assert (lens >= 5).all()
            continue
        np.testing.assert_array_equal(ids, ref[0], err_msg=f"cluster {cluster} group {group}")
        np.testing.assert_array_equal(lens, ref[1])
        np.testing.assert_array_equal(lg, ref[2])
