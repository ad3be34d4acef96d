"""Host-side C++ mirror of whisper.h (csrc/host) against the reference's OWN implementation of the same
functions (oracle/_ref: whisper.cpp:519-665 Reader/decode/language table, wav_util.cpp:18-129, mmap_file.cpp)
compiled in place, plus pinned expectations that hold without it. No device work here."""
import os
import struct
import subprocess

import numpy as np
import pytest

from whisper_tflite_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host():
    from whisper_tflite_b200 import build, host
    build.build_all()
    host.load()
    return host


@pytest.fixture(scope="module")
def ref(oracle_built):
    if not oracle_built.have_ref():
        pytest.skip("oracle/_ref not built (reference tree absent)")
    return oracle_built


def make_blob(tmp_path, n_tokens=50257, name="blob.bin", tokens=None):
    toks = tokens if tokens is not None else synth.synth_vocab_tokens(n_tokens)
    p = str(tmp_path / name)
    synth.write_vocab_blob(p, synth.mel_filters(), toks)
    return open(p, "rb").read()


def test_exports(host):
    lib = host.load()
    for s in host.EXPORTED:
        assert getattr(lib, s) is not None


def test_jni_bridge_compiles_and_exports_the_four_natives(host):
    obj = os.path.join(host.HOST_DIR, "jni_whisper.o")
    assert os.path.exists(obj)
    syms = subprocess.run(["nm", "-g", "--defined-only", obj], capture_output=True, text=True).stdout
    for fn in ("create", "destroy", "transcribeBuffer", "transcribeFile"):  # bindings/java/whisper.tflite.cpp:17-71
        assert f"Java_io_github_jerinphilip_whisper_EngineNative_{fn}" in syms


@pytest.mark.parametrize("multilingual", [True, False])
def test_reader_matches_reference(host, ref, tmp_path, multilingual):
    blob = make_blob(tmp_path, 50257 if multilingual else 50256)
    mine = host.Blob(blob, multilingual)
    theirs = ref.RefVocab(blob, multilingual)
    assert mine.n_vocab == theirs.n_vocab
    assert mine.special == theirs.special
    np.testing.assert_array_equal(mine.filters, theirs.filters)
    n_total = 51864 + int(multilingual)
    assert mine.n_entries == n_total
    for i in list(range(0, 50257, 997)) + list(range(50250, n_total)):
        assert mine.token(i) == theirs.token(i), i
    assert mine.token(n_total) is None and theirs.token(n_total) is None


def test_reader_pinned_names(host, tmp_path):
    """Expectations read off whisper.cpp:577-604 / whisper.h:66-91; hold without oracle/_ref."""
    b = host.Blob(make_blob(tmp_path), True)
    assert b.n_vocab == 51865
    assert b.special == {"eot": 50257, "sot": 50258, "translate": 50358, "transcribe": 50359, "prev": 50361,
                         "solm": 50362, "not": 50363, "beg": 50364}
    assert b.token(17) == b" t17"
    assert b.token(50257) == b"<|endoftranscript|>"
    assert b.token(50258) == b"<|startoftranscript_|>"
    assert b.token(50259) == b"<|lang-en|>" and b.token(50261) == b"<|lang-de|>"
    assert b.token(50358) == b"<|translate|>" and b.token(50359) == b"<|transcribe|>"
    assert b.token(50360) == b"<|e50360|>"
    assert b.token(50361) == b"<|PREV|>" and b.token(50363) == b"<|notimestamps|>"
    assert b.token(50364) == b"<|timestampbegin|>" and b.token(50365) == b"<|TT1|>"
    e = host.Blob(make_blob(tmp_path, 50256, "en.bin"), False)
    assert e.n_vocab == 50256 and e.special["eot"] == 50256 and e.special["not"] == 50362
    assert e.token(50256) == b"<|endoftranscript|>" and e.token(51863) == b"<|TT1500|>"


def test_reader_token_with_embedded_nul_and_max_len(host, ref, tmp_path):
    toks = synth.synth_vocab_tokens(300)
    toks[5] = b"ab\0cd"          # the reference copies through a C string: truncated at the NUL (whisper.cpp:568-571)
    toks[6] = b"x" * 255         # longest token the reference's 256-byte buffer holds
    toks[7] = b""
    blob = make_blob(tmp_path, tokens=toks)
    mine, theirs = host.Blob(blob, True), ref.RefVocab(blob, True)
    for i in (4, 5, 6, 7, 8, 299, 300, 50257):
        assert mine.token(i) == theirs.token(i), i
    assert mine.token(5) == b"ab" and mine.token(6) == b"x" * 255 and mine.token(7) == b""


def test_decode_matches_reference(host, ref, tmp_path):
    blob = make_blob(tmp_path)
    mine, theirs = host.Blob(blob, True), ref.RefVocab(blob, True)
    rs = np.random.RandomState(0)
    cases = [[50258, 50261, 50359, 50363, 11, 22, 33, 50257, 44, 55],   # stops after EOT
             [50258, 50261, 50359, 50363] + [7] * 27,                  # degenerate repeat, no EOT
             [50257], [], [0], [51864, 50364, 50365]]
    for _ in range(20):
        n = rs.randint(1, 31)
        ids = list(rs.randint(0, 51865, size=n))
        if rs.rand() < 0.5:
            ids.insert(rs.randint(0, n), 50257)
        cases.append(ids)
    for ids in cases:
        for omit in (False, True):
            assert mine.decode(ids, omit) == theirs.decode(ids, omit), (ids, omit)
            assert mine.decode(ids, omit, i32=True) == theirs.decode(ids, omit)
    assert mine.decode([50258, 50261, 50359, 50363, 11, 50257, 12]) == \
        b"<|startoftranscript_|><|lang-de|><|transcribe|><|notimestamps|> t11<|endoftranscript|>"
    assert mine.decode([50258, 50261, 50359, 50363, 11, 50257, 12], True) == b" t11"


def test_language_table(host, ref):
    assert ref.ref_language_count() == 100
    for i in range(100):
        code = ref.ref_lang_code(i)
        assert host.lang_code(i) == code
        assert host.language_id(code) == ref.ref_language_id(code) == i
    assert host.language_id("de") == 2  # -> prompt token 50261 (whisper.cpp:327)
    assert host.language_id("xx") == ref.ref_language_id("xx") == 100


def test_remove_extra_spaces(host, ref):
    rs = np.random.RandomState(1)
    cases = ["", " ", "  ", "a  b   c", "  lead", "trail   ", "no_spaces", " a b ", "tab\t \t  x"]
    for _ in range(50):
        cases.append("".join(rs.choice(list("ab  \t"), size=rs.randint(0, 40))))
    for s in cases:
        assert host.remove_extra_spaces(s) == ref.ref_remove_extra_spaces(s), repr(s)
    assert host.remove_extra_spaces("a   b  c ") == "a b c "


def write_wav(path, samples, fmt=1, channels=1, rate=16000, extra_chunk=False, riff_size=None):
    """Canonical RIFF/WAVE file: 'fmt ' (16 bytes) [+ LIST chunk] + 'data'."""
    if fmt == 1:
        payload = np.asarray(samples, dtype="<i2").tobytes()
        bits = 16
    else:
        payload = np.asarray(samples, dtype="<f4").tobytes()
        bits = 32
    block = channels * bits // 8
    body = b"WAVE" + b"fmt " + struct.pack("<IHHIIHH", 16, fmt, channels, rate, rate * block, block, bits)
    if extra_chunk:
        body += b"LIST" + struct.pack("<I", 6) + b"abcdef"
    body += b"data" + struct.pack("<I", len(payload)) + payload
    size = len(body) if riff_size is None else riff_size
    with open(path, "wb") as f:
        f.write(b"RIFF" + struct.pack("<I", size) + body)


def test_wav_read_legacy_matches_reference(host, ref, tmp_path):
    rs = np.random.RandomState(2)
    pcm = rs.randint(-32768, 32767, size=16000).astype(np.int16)
    p = str(tmp_path / "a.wav")
    write_wav(p, pcm)
    mine, theirs = host.wav_read_legacy(p), ref.ref_wav_read_legacy(p)
    np.testing.assert_array_equal(mine, theirs)
    # quirks kept (wav_util.cpp:29, :61-69): the 8-byte data-chunk header is decoded as 4 leading samples, the RIFF
    # size (file size - 8 = 4 + 24 + 8 + 32000) is taken as the payload size -> 14 extra zero samples at the end
    assert len(mine) == (4 + 24 + 8 + 32000) // 2
    np.testing.assert_array_equal(mine[4:4 + 16000], pcm.astype(np.float32) / np.float32(32767))
    assert np.all(mine[4 + 16000:] == 0)
    # IEEE float payload
    pf = str(tmp_path / "f.wav")
    write_wav(pf, rs.uniform(-1, 1, 4000).astype(np.float32), fmt=3)
    np.testing.assert_array_equal(host.wav_read_legacy(pf), ref.ref_wav_read_legacy(pf))
    # truncated file: RIFF size says more than the file holds -> tail stays zero
    pt = str(tmp_path / "t.wav")
    write_wav(pt, pcm[:1000], riff_size=4 + 24 + 8 + 32000)
    np.testing.assert_array_equal(host.wav_read_legacy(pt), ref.ref_wav_read_legacy(pt))
    # not a WAV / missing -> empty vector
    pb = str(tmp_path / "bad.wav")
    open(pb, "wb").write(b"RIFX" + b"\0" * 100)
    assert len(host.wav_read_legacy(pb)) == len(ref.ref_wav_read_legacy(pb)) == 0
    assert len(host.wav_read_legacy(str(tmp_path / "missing.wav"))) == 0


def test_wav_read_matches_reference(host, ref, tmp_path):
    rs = np.random.RandomState(3)
    mono = rs.randint(-32768, 32767, size=5000).astype(np.int16)
    p = str(tmp_path / "m.wav")
    write_wav(p, mono, extra_chunk=True)
    np.testing.assert_array_equal(host.wav_read(p), ref.ref_wav_read(p))
    np.testing.assert_array_equal(host.wav_read(p), mono.astype(np.float32) / np.float32(32767))
    st = rs.randint(-32768, 32767, size=2 * 3000).astype(np.int16)
    ps = str(tmp_path / "s.wav")
    write_wav(ps, st, channels=2)
    np.testing.assert_array_equal(host.wav_read(ps), ref.ref_wav_read(ps))


def test_mmap_file(host, tmp_path):
    p = str(tmp_path / "blob.bin")
    data = bytes(range(200))
    open(p, "wb").write(data)
    size, first = host.mmap_probe(p)
    assert size == 200 and first == data[:16]
    with pytest.raises(RuntimeError, match="Failed to open file"):  # mmap_file.cpp:15-17
        host.mmap_probe(str(tmp_path / "missing.bin"))


def test_create_engine_error_paths_without_device(host, tmp_path):
    """Unknown engine type -> nullptr (whisper.cpp:785-789); missing vocab -> MmapFile's runtime_error."""
    assert not host.load().wth_engine_create(7, b"/nonexistent/prefix", b"/nonexistent/vocab", 1)
    with pytest.raises(RuntimeError, match="Failed to open file"):
        host.HostEngine(host.ENGINE_ENCDEC, "/nonexistent/prefix", str(tmp_path / "missing.bin"), True)
