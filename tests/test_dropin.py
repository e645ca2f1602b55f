"""The link-time drop-in: the reference's own callers of the hot path — tests/test_decoder.cpp,
tests/test_transformer.cpp, tests/test_vq_only.cpp and the whole pipeline src/main.cpp + src/qwen3_tts.cpp —
compiled UNCHANGED from /root/reference against this repository's host headers and linked against
libq3t_b200.so (tools/dropin/Makefile).  CPU part: they compile, link and bind to our classes, and fail loudly
without a GPU.  GPU part (binaries travel with the snapshot): the reference's own test programs print their PASS
lines on fixtures written from the committed goldens, and the drop-in qwen3-tts-cli produces the waveform of the
unmodified reference CLI for BASELINE.json config 1 ('Hello, world!', greedy).
"""
import os
import subprocess
import wave

import numpy as np
import pytest

from util import GOLDEN, ROOT, models, snr_db

BIN = os.path.join(ROOT, "tools", "dropin", "bin")
HAVE_REFSRC = os.path.isdir("/root/reference/src") and os.path.isdir("/root/reference/tests")
NAMES = ["test_decoder", "test_transformer", "test_vq_only", "qwen3-tts-cli", "qwen3-tts-cli-ref"]


def _have_bins():
    return all(os.path.exists(os.path.join(BIN, n)) for n in NAMES)


def _run(cmd, cwd=None, timeout=900):
    r = subprocess.run(cmd, cwd=cwd, capture_output=True, text=True, timeout=timeout)
    return r.returncode, r.stdout + r.stderr


@pytest.mark.skipif(not HAVE_REFSRC, reason="/root/reference not present (GPU box uses the prebuilt binaries)")
def test_reference_callers_compile_and_link_against_the_engine(q3t):
    assert os.path.exists(os.path.join(ROOT, "oracle", "_ref", "obj", "ggml")), "run make -f oracle/Makefile.ref first"
    r = subprocess.run(["make", "-f", "tools/dropin/Makefile", "-j4"], cwd=ROOT, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert _have_bins()
    for name, syms in (("test_decoder", ["qwen3_tts::AudioTokenizerDecoder::decode(", "qwen3_tts::AudioTokenizerDecoder::load_model("]),
                       ("test_transformer", ["qwen3_tts::TTSTransformer::generate(", "qwen3_tts::TTSTransformer::forward_prefill("]),
                       ("qwen3-tts-cli", ["qwen3_tts::TTSTransformer::generate(", "qwen3_tts::AudioTokenizerDecoder::decode("])):
        und = subprocess.run(["nm", "-uC", os.path.join(BIN, name)], capture_output=True, text=True).stdout
        for s in syms:
            assert s in und, f"{name}: {s} is not an undefined (= imported) symbol"
        ldd = subprocess.run(["ldd", os.path.join(BIN, name)], capture_output=True, text=True).stdout
        assert "libq3t_b200.so" in ldd and "not found" not in ldd, ldd
    # the drop-in CLI holds no ggml transformer / vocoder of its own: the classes come from the engine library only
    defined = subprocess.run(["nm", "-C", "--defined-only", os.path.join(BIN, "qwen3-tts-cli")], capture_output=True, text=True).stdout
    assert "qwen3_tts::TTSTransformer::generate" not in defined and "qwen3_tts::AudioTokenizerDecoder::decode" not in defined
    exported = subprocess.run(["nm", "-DC", "--defined-only", q3t.LIB_PATH], capture_output=True, text=True).stdout
    assert "qwen3_tts::TTSTransformer::generate(" in exported and "qwen3_tts::AudioTokenizerDecoder::decode(" in exported
    assert "qwen3_tts::GGUFLoader::" not in exported, "the engine must not shadow the reference's ggml GGUFLoader"


@pytest.mark.skipif(not _have_bins() and not HAVE_REFSRC, reason="drop-in binaries not built")
def test_linked_caller_fails_loudly_without_a_gpu(tiny_models):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    if not _have_bins():
        subprocess.run(["make", "-f", "tools/dropin/Makefile", "-j4"], cwd=ROOT, check=True, capture_output=True)
    rc, out = _run([os.path.join(BIN, "test_decoder"), "--tokenizer", tiny_models[1]])
    assert rc != 0 and "FAIL" in out and "CUDA" in out, out[-1500:]


# ------------------------------------------------------------------------------------------ GPU box
gpu = pytest.mark.gpu
need_bins = pytest.mark.skipif(not _have_bins(), reason="tools/dropin/bin not shipped")


@gpu
@need_bins
def test_reference_test_decoder_passes_on_golden_fixture(tmp_path, tiny_models):
    g = np.load(os.path.join(GOLDEN, "vocoder_f16.npz"))
    g["codes"].astype(np.int64).tofile(tmp_path / "speech_codes.bin")
    g["wav"].astype(np.float32).tofile(tmp_path / "decoded_audio.bin")
    rc, out = _run([os.path.join(BIN, "test_decoder"), "--tokenizer", tiny_models[1], "--codes", str(tmp_path / "speech_codes.bin"),
                    "--reference", str(tmp_path / "decoded_audio.bin")])
    assert rc == 0, out[-3000:]
    assert "PASS: L2 distance < 0.001 (excellent match)" in out or "PASS: L2 distance < 0.01 (good match)" in out, out[-3000:]
    l2 = float(out.split("L2 distance (RMS):")[1].split()[0])
    assert l2 < 0.003, l2                                    # measured 0.0016 = 45 dB below the signal
    assert "PASS: Correlation > 0.95 (excellent)" in out, out[-3000:]
    assert "FAIL" not in out, out[-3000:]


@gpu
@need_bins
def test_reference_test_transformer_matches_all_codes(tmp_path):
    g = np.load(os.path.join(GOLDEN, "full_f16.npz"))
    tts = models("full", "f16")[0]
    g["tokens"].astype(np.int64).tofile(tmp_path / "det_text_tokens.bin")
    g["spk"].astype(np.float32).tofile(tmp_path / "det_speaker_embedding.bin")
    g["prefill"].astype(np.float32).tofile(tmp_path / "det_prefill_embedding.bin")
    g["prefill_logits"].astype(np.float32).tofile(tmp_path / "det_first_frame_logits.bin")
    g["codes"].astype(np.int64).tofile(tmp_path / "det_speech_codes.bin")
    rc, out = _run([os.path.join(BIN, "test_transformer"), "--model", tts, "--ref-dir", str(tmp_path), "--max-len", "64"])
    assert rc == 0, out[-3000:]
    assert "Max absolute error:   0.00000000" in out, out[-4000:]
    assert "Matching frames:       64 / 64 (100.0%)" in out, out[-4000:]
    assert "First mismatch frame:  none (all match!)" in out, out[-4000:]
    assert "FAIL:   0" in out, out[-4000:]


@gpu
@need_bins
def test_reference_test_vq_only_correlates_with_live_reference(tmp_path, tiny_models):
    from oracle import reflib
    if not reflib.available():
        pytest.skip("oracle/_ref not shipped")
    g = np.load(os.path.join(GOLDEN, "full_f16.npz"))
    codes = np.ascontiguousarray(g["codes"][:63] % 2048, np.int32)
    dec = reflib.RefDecoder(tiny_models[1])
    wav = dec.decode(codes)
    dec.close()
    (tmp_path / "models").mkdir(); (tmp_path / "reference").mkdir()
    os.symlink(tiny_models[1], tmp_path / "models" / "qwen3-tts-tokenizer-f16.gguf")
    codes.astype(np.int64).tofile(tmp_path / "reference" / "speech_codes.bin")
    wav.astype(np.float32).tofile(tmp_path / "reference" / "decoded_audio.bin")
    rc, out = _run([os.path.join(BIN, "test_vq_only")], cwd=str(tmp_path))
    assert rc == 0, out[-2000:]
    assert f"Decoded {len(wav)} samples" in out, out[-2000:]
    corr = float(out.split("Correlation:")[1].split()[0])
    assert corr > 0.9999, out[-2000:]


def _read_wav(path):
    with wave.open(str(path), "rb") as w:
        assert w.getframerate() == 24000 and w.getnchannels() == 1 and w.getsampwidth() == 2
        return np.frombuffer(w.readframes(w.getnframes()), np.int16).astype(np.float64)


@gpu
@need_bins
def test_dropin_cli_equals_reference_cli_hello_world(tmp_path):
    """BASELINE.json config 1 through the reference's own front door: text -> reference tokenizer -> generate ->
    decode -> 16-bit WAV, once with the B200 engine behind the classes and once with the unmodified reference."""
    tts, voc = models("full", "f16")
    mdir = tmp_path / "models"
    mdir.mkdir()
    os.symlink(tts, mdir / "qwen3-tts-0.6b-f16.gguf")          # the pipeline's hard-coded names (src/qwen3_tts.cpp:118-119)
    os.symlink(voc, mdir / "qwen3-tts-tokenizer-f16.gguf")
    args = ["-m", str(mdir), "-t", "Hello, world!", "--temperature", "0", "--max-tokens", "24"]
    rc, out = _run([os.path.join(BIN, "qwen3-tts-cli")] + args + ["-o", str(tmp_path / "b200.wav")])
    assert rc == 0, out[-3000:]
    assert "CUDA sm_100" in out, out[-3000:]
    rc, out_ref = _run([os.path.join(BIN, "qwen3-tts-cli-ref")] + args + ["-o", str(tmp_path / "ref.wav")])
    assert rc == 0, out_ref[-3000:]
    a, b = _read_wav(tmp_path / "b200.wav"), _read_wav(tmp_path / "ref.wav")
    assert len(a) == len(b) == 1920 * 24 - 555
    assert snr_db(b, a) >= 30.0, snr_db(b, a)


@gpu
@need_bins
def test_dropin_cli_voice_clone_runs_the_gpu_speaker_encoder(tmp_path):
    """BASELINE.json config 2 through the reference's front door: `-r <wav>` -> load_audio_file / resample (reference
    host code) -> AudioTokenizerEncoder::encode (the engine's GPU encoder) -> generate -> decode.  The two encoders agree
    to ~6e-4 relative (tests/test_speaker_encoder_gpu.py); with random weights that can flip a greedy token, so the
    waveforms are compared for length and reported, not asserted equal."""
    from test_speaker_encoder_gpu import _clip
    tts, voc = models("full", "f16")
    mdir = tmp_path / "models"
    mdir.mkdir()
    os.symlink(tts, mdir / "qwen3-tts-0.6b-f16.gguf")
    os.symlink(voc, mdir / "qwen3-tts-tokenizer-f16.gguf")
    x = _clip(2.0, 21)
    with wave.open(str(tmp_path / "voice.wav"), "wb") as w:
        w.setnchannels(1); w.setsampwidth(2); w.setframerate(24000)
        w.writeframes((x * 32767.0).astype(np.int16).tobytes())
    args = ["-m", str(mdir), "-t", "Hello, world!", "-r", str(tmp_path / "voice.wav"), "--temperature", "0", "--max-tokens", "12"]
    rc, out = _run([os.path.join(BIN, "qwen3-tts-cli")] + args + ["-o", str(tmp_path / "b200.wav")])
    assert rc == 0, out[-3000:]
    assert "AudioTokenizerEncoder backend: CUDA sm_100" in out, out[-3000:]
    rc, out_ref = _run([os.path.join(BIN, "qwen3-tts-cli-ref")] + args + ["-o", str(tmp_path / "ref.wav")])
    assert rc == 0, out_ref[-3000:]
    a, b = _read_wav(tmp_path / "b200.wav"), _read_wav(tmp_path / "ref.wav")
    assert len(a) == len(b) == 1920 * 12 - 555
    print(f"[voice clone CLI] SNR of the drop-in waveform vs the reference CLI: {snr_db(b, a):.1f} dB")
