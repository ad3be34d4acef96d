"""Generates tests/golden/model_golden_<size>.npz — run IN THE BUILD CONTAINER (needs transformers).

Independent implementation used as the pin: HF `transformers` WhisperForConditionalGeneration
(modeling_whisper.py) loaded with the seeded weights of `whisper_tflite_b200.synth.init_weights`,
driven by the reference's greedy loop semantics (whisper.cpp:334-400: full re-forward per step,
last-index argmax, EOT stop, <=27 new tokens). Inputs are regenerated from seeds by the tests, so
only small output samples are stored: a strided sample of the encoder states, the top-8 logits of
every step, and the token ids.

    python tests/golden/make_model_golden.py micro tiny
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import mel as omel  # noqa: E402
from oracle import model_ref  # noqa: E402
from whisper_tflite_b200 import convert, synth  # noqa: E402


def hf_model(dims, W):
    from transformers import WhisperConfig, WhisperForConditionalGeneration
    cfg = WhisperConfig(vocab_size=dims.n_vocab, num_mel_bins=dims.n_mels, d_model=dims.d_model,
                        encoder_layers=dims.n_enc_layers, decoder_layers=dims.n_dec_layers,
                        encoder_attention_heads=dims.n_heads, decoder_attention_heads=dims.n_heads,
                        encoder_ffn_dim=4 * dims.d_model, decoder_ffn_dim=4 * dims.d_model,
                        max_source_positions=dims.n_audio_ctx, max_target_positions=dims.n_text_ctx,
                        attn_implementation="eager")
    m = WhisperForConditionalGeneration(cfg).eval()
    sd = {k: torch.from_numpy(np.array(v)) for k, v in convert.canonical_to_hf(W).items()}
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all("proj_out" in k for k in missing), (missing, unexpected)
    return m


def main(sizes, n_clips=2, seed=0):
    torch.set_grad_enabled(False)
    F = synth.mel_filters()
    pcm = synth.synth_pcm(n_clips, seed=seed)
    mels = np.stack([omel.ref_log_mel(p, F, 0) for p in pcm])
    for size in sizes:
        dims = synth.DIMS[size]
        W = synth.init_weights(dims, seed=seed)
        m = hf_model(dims, W)
        prompt0 = model_ref.reference_prompt(True)
        eot = 50257
        enc_samples, all_ids, all_top_idx, all_top_val, all_lens = [], [], [], [], []
        for c in range(n_clips):
            xa = m.model.encoder(torch.from_numpy(mels[c:c + 1])).last_hidden_state  # [1,1500,d]
            prompt = list(prompt0)
            top_idx = np.zeros((27, 8), np.int64)
            top_val = np.zeros((27, 8), np.float32)
            for s, _i in enumerate(range(len(prompt0) - 1, 30)):
                dec = m.model.decoder(input_ids=torch.tensor([prompt]), encoder_hidden_states=xa).last_hidden_state
                row = m.proj_out(dec)[0, -1].numpy()
                nxt = model_ref.argmax_last(row)
                order = np.argsort(-row, kind="stable")[:8]
                top_idx[s], top_val[s] = order, row[order]
                prompt.append(nxt)
                if nxt == eot:
                    break
            ids = np.full(31, -1, np.int64)
            ids[:len(prompt)] = prompt
            all_ids.append(ids)
            all_lens.append(len(prompt))
            all_top_idx.append(top_idx)
            all_top_val.append(top_val)
            enc_samples.append(xa[0].numpy()[::25, ::(dims.d_model // 32)].copy())  # [60, 32]
            # independent check of the numpy oracle while we are here
            xo = model_ref.encoder_forward(W, dims, mels[c])
            err = np.abs(xo - xa[0].numpy()).max() / np.abs(xa[0].numpy()).max()
            ids_o = model_ref.greedy_decode(W, dims, xo, prompt0, eot)
            print(f"{size} clip{c}: enc rel-to-max err oracle vs HF {err:.2e}; tokens equal: {ids_o == prompt}; "
                  f"len {len(prompt)}; min top1-top2 margin {np.min(top_val[:len(prompt) - 4, 0] - top_val[:len(prompt) - 4, 1]):.3e}")
        out = os.path.join(ROOT, "tests", "golden", f"model_golden_{size}.npz")
        np.savez_compressed(out, seed=seed, n_clips=n_clips, enc_sample=np.stack(enc_samples),
                            ids=np.stack(all_ids), lens=np.array(all_lens), top_idx=np.stack(all_top_idx),
                            top_val=np.stack(all_top_val), prompt=np.array(prompt0), eot=eot)
        print("wrote", out, os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main(sys.argv[1:] or ["micro", "tiny"])
