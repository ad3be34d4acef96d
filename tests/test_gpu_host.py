"""The reference-facing C++ API (csrc/host mirror of whisper.h, driven as app/encdec and the JNI bridge drive
it) on the device, against the CPU oracle on the same synthetic PCM, seeded weights and vocab blob:
log_mel_spectrogram within 1e-4 x range; EncDec / Monolith transcripts equal to decode(oracle greedy ids) in
the fp32 verification mode (token identity => string identity)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from oracle import model_ref
from tests.gpu_util import weight_file
from tests.test_host_cpu import write_wav
from whisper_tflite_b200 import synth

pytestmark = pytest.mark.gpu
EOT = 50257


@pytest.fixture(scope="module")
def assets(tmp_path_factory):
    """micro-model weights as <prefix>.b200w + multilingual and English vocab blobs."""
    d = tmp_path_factory.mktemp("host_assets")
    path, dims, W = weight_file("micro")
    prefix = str(d / "whisper-micro")
    os.symlink(path, prefix + ".b200w")
    vm, ve = str(d / "filters_vocab_multilingual.bin"), str(d / "filters_vocab_en.bin")
    synth.write_vocab_blob(vm, synth.mel_filters(), synth.synth_vocab_tokens(50257))
    synth.write_vocab_blob(ve, synth.mel_filters(), synth.synth_vocab_tokens(50256))
    return {"prefix": prefix, "vocab_multi": vm, "vocab_en": ve, "dims": dims, "W": W, "dir": str(d)}


@pytest.fixture(scope="module")
def host():
    from whisper_tflite_b200 import build, host
    build.build_all()
    os.environ["WHISPER_B200_PRECISION"] = "fp32"  # token identity is promised in the verification mode
    os.environ["WHISPER_B200_MAX_BATCH"] = "2"     # forces transcribe_batch to chunk
    host.load()
    return host


def oracle_ids(assets, pcm, oracle_built, filters, prompt, eot):
    mel = oracle_built.oracle_log_mel(pcm, filters, 8)
    xa = model_ref.encoder_forward(assets["W"], assets["dims"], mel)
    return model_ref.greedy_decode(assets["W"], assets["dims"], xa, list(prompt), eot)


def test_log_mel_spectrogram_free_function(host, oracle_built, filters):
    pcm = synth.synth_pcm(2, seed=5)
    for c, n in ((0, 480000), (1, 160 * 211 + 57)):  # full window, ragged short clip (n_len = n / 160)
        x = pcm[c][:n]
        got = host.log_mel_spectrogram(x, filters)
        ref = oracle_built.oracle_log_mel(x, filters, 8)
        assert got.shape == ref.shape == (80, n // 160)
        assert np.abs(got - ref).max() <= 1e-4 * (ref.max() - ref.min())
    # a different filterbank re-arms the frontend (dense rows, 40 bins)
    dense = np.abs(np.random.RandomState(0).standard_normal((40, 201))).astype(np.float32) * 0.01
    got = host.log_mel_spectrogram(pcm[0], dense)
    ref = oracle_built.oracle_log_mel(pcm[0], dense, 8)
    assert np.abs(got - ref).max() <= 1e-4 * (ref.max() - ref.min())


def test_encdec_transcribe_matches_oracle(host, assets, oracle_built, filters):
    eng = host.HostEngine(host.ENGINE_ENCDEC, assets["prefix"], assets["vocab_multi"], True)
    blob = host.Blob(open(assets["vocab_multi"], "rb").read(), True)
    pcm = synth.synth_pcm(3, seed=11)
    texts = []
    for c in range(3):
        n = 480000 if c != 1 else 300000  # short input is zero-padded to 30 s (whisper.cpp:753)
        text, size_after = eng.transcribe(pcm[c][:n])
        assert size_after == 480000
        padded = np.zeros(480000, np.float32)
        padded[:n] = pcm[c][:n]
        ids = oracle_ids(assets, padded, oracle_built, filters, (50258, 50261, 50359, 50363), EOT)
        assert text == blob.decode(ids, False)
        assert text.startswith(b"<|startoftranscript_|><|lang-de|><|transcribe|><|notimestamps|>")
        texts.append(text)
    # batch of 3 through a max_batch = 2 engine == the single-clip calls
    batch = pcm.copy()
    batch[1, 300000:] = 0
    assert eng.transcribe_batch(batch) == texts
    eng.close()


def test_monolith_prompts(host, assets, oracle_built, filters):
    pcm = synth.synth_pcm(1, seed=12)[0]
    multi = host.HostEngine(host.ENGINE_MONOLITH, assets["prefix"], assets["vocab_multi"], True)
    text, _ = multi.transcribe(pcm)
    blob = host.Blob(open(assets["vocab_multi"], "rb").read(), True)
    assert text == blob.decode(oracle_ids(assets, pcm, oracle_built, filters, (50258, 50261, 50359, 50363), EOT))
    multi.close()
    # English vocabulary (app/minimal: multilingual = false): prompt [sot, notimestamps] = kGoldenGeneratedIDs[:2]
    en = host.HostEngine(host.ENGINE_MONOLITH, assets["prefix"], assets["vocab_en"], False)
    text, _ = en.transcribe(pcm)
    blob_en = host.Blob(open(assets["vocab_en"], "rb").read(), False)
    ids = oracle_ids(assets, pcm, oracle_built, filters, (50257, 50362), 50256)
    assert ids[:2] == [50257, 50362]
    assert text == blob_en.decode(ids)
    en.close()


def test_encdec_cli_on_wav_file(host, assets, oracle_built, filters):
    """app/encdec end to end: WAV file -> wav_read_legacy (quirks included) -> transcript on stdout."""
    rs = np.random.RandomState(4)
    pcm16 = (synth.synth_pcm(1, seed=13)[0][:200000] * 20000).astype(np.int16)
    wav = os.path.join(assets["dir"], "clip.wav")
    write_wav(wav, pcm16)
    env = dict(os.environ, WHISPER_B200_PRECISION="fp32", WHISPER_B200_MAX_BATCH="1")
    out = subprocess.run([host.ENCDEC_EXE, "--model-prefix", assets["prefix"], "--vocab", assets["vocab_multi"],
                          "--input", wav], capture_output=True, env=env, timeout=300)
    assert out.returncode == 0, out.stderr.decode()
    samples = host.wav_read_legacy(wav)
    padded = np.zeros(480000, np.float32)
    padded[:len(samples)] = samples
    ids = oracle_ids(assets, padded, oracle_built, filters, (50258, 50261, 50359, 50363), EOT)
    blob = host.Blob(open(assets["vocab_multi"], "rb").read(), True)
    assert out.stdout.rstrip(b"\n").split(b"\n")[-1] == blob.decode(ids)
    # usage error: missing flags -> non-zero exit, nothing transcribed
    bad = subprocess.run([host.ENCDEC_EXE, "--vocab", assets["vocab_multi"]], capture_output=True, timeout=60)
    assert bad.returncode != 0
    # app/minimal prints the space-collapsed transcript
    out2 = subprocess.run([host.MINIMAL_EXE, assets["prefix"], assets["vocab_en"], wav], capture_output=True, env=env,
                          timeout=300)
    assert out2.returncode == 0, out2.stderr.decode()
    ids2 = oracle_ids(assets, padded, oracle_built, filters, (50257, 50362), 50256)
    blob_en = host.Blob(open(assets["vocab_en"], "rb").read(), False)
    expect = host.remove_extra_spaces(blob_en.decode(ids2).decode()).encode()
    assert expect in out2.stdout


def test_language_task_and_special_token_options(host, assets, oracle_built, filters):
    """what the reference hard-codes — language "de" (whisper.cpp:327), transcribe
    (:336), omit_special_tokens = false (:733, :766) — as options whose defaults are the reference's. The prompt the
    engine builds must be [sot, 50259 + language_id(code), transcribe | translate, notimestamps]: transcripts equal
    decode(oracle greedy ids for that prompt) in the verification mode."""
    pcm = synth.synth_pcm(1, seed=14)[0]
    blob = host.Blob(open(assets["vocab_multi"], "rb").read(), True)
    for etype in (host.ENGINE_ENCDEC, host.ENGINE_MONOLITH):
        eng = host.HostEngine(etype, assets["prefix"], assets["vocab_multi"], True)
        default_text, _ = eng.transcribe(pcm)
        ids_de = oracle_ids(assets, pcm, oracle_built, filters, (50258, 50261, 50359, 50363), EOT)
        assert default_text == blob.decode(ids_de)                        # defaults: "de", transcribe, specials kept
        assert not eng.set_language("xx")                                 # unknown code: refused, nothing changes
        assert eng.transcribe(pcm)[0] == default_text
        assert eng.set_language("fr") and host.language_id("fr") == 6
        ids_fr = oracle_ids(assets, pcm, oracle_built, filters, (50258, 50259 + 6, 50359, 50363), EOT)
        text_fr, _ = eng.transcribe(pcm)
        assert text_fr == blob.decode(ids_fr) and text_fr.startswith(b"<|startoftranscript_|><|lang-fr|><|transcribe|>")
        eng.set_task(True)                                                # translate: token 50358 (unshifted, whisper.cpp:218-226)
        ids_tr = oracle_ids(assets, pcm, oracle_built, filters, (50258, 50259 + 6, 50358, 50363), EOT)
        assert eng.transcribe(pcm)[0] == blob.decode(ids_tr)
        eng.set_omit_special_tokens(True)                                 # decode keeps only ids < token_eot (whisper.cpp:641-650)
        assert eng.transcribe(pcm)[0] == blob.decode(ids_tr, True)
        assert b"<|" not in blob.decode(ids_tr, True)
        eng.close()


def test_int16_ingest_matches_host_conversion(host, assets):
    """16-bit samples go to the device as they are; the `/ INT16_MAX` of wav_read_legacy /
    wav_read (wav_util.cpp:70-73, :120-122) happens in the mel kernel's staging load. Transcripts must equal those of
    the f32 call on host-converted samples (bit-equal mel => identical ids), for both engine kinds, through a
    max_batch = 2 engine that has to chunk 3 clips."""
    pcm16 = np.clip(synth.synth_pcm(3, seed=15) * 25000.0, -32768, 32767).astype(np.int16)
    pcm16[1, 300000:] = 0
    as_f32 = (pcm16.astype(np.float32) / np.float32(32767.0)).astype(np.float32)
    for etype in (host.ENGINE_ENCDEC, host.ENGINE_MONOLITH):
        eng = host.HostEngine(etype, assets["prefix"], assets["vocab_multi"], True)
        assert eng.transcribe_batch_i16(pcm16) == eng.transcribe_batch(as_f32)
        # short clips are zero-padded to 30 s like the f32 path (whisper.cpp:753)
        assert eng.transcribe_batch_i16(pcm16[:1, :200000])[0] == eng.transcribe_batch(as_f32[:1, :200000])[0]
        eng.close()
