#!/usr/bin/env python3
"""TEST INFRASTRUCTURE — writes the golden vectors under tests/golden/ by running the UNMODIFIED
reference (oracle/_ref/libq3t_ref.so, built by oracle/Makefile.ref from /root/reference) on the
seeded random-init GGUFs of tools/make_random_gguf.py.

    make -f oracle/Makefile.ref -j8 && python oracle/gen_golden.py [--full]

The reference build pins AVX2+FMA+F16C, so its F32 summation order — and every vector written
here — is identical on any x86-64 host with those extensions (checked: build container vs GPU box).
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import reflib  # noqa: E402
from util import models, HELLO_TINY, HELLO_FULL, GOLDEN  # noqa: E402


def golden_inputs(seed=0):
    rng = np.random.default_rng(seed)
    spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
    step_x = (rng.standard_normal(1024) * 0.05).astype(np.float32)
    return spk, step_x


def gen_tts(preset, wtype, tokens, n_frames, out):
    tts, _ = models(preset, wtype)
    ref = reflib.RefTTS(tts)
    spk, step_x = golden_inputs()
    pre, trail, pad = ref.build_prompt(tokens, spk)
    pre_ns, trail_ns, _ = ref.build_prompt(tokens, None, -1)           # no speaker, no language row
    ref.init_kv(64)
    hid, logits = ref.prefill(pre)
    step_logits, step_hidden = ref.step(step_x, pre.shape[0])
    cb0 = int(step_logits[:2048].argmax())
    cp_codes = ref.predict_codes_ar(step_hidden, cb0, 0.0, 0)
    codes = ref.generate(tokens, spk, n_frames, rep_penalty=1.05, temperature=0.0)
    codes_nospk = ref.generate(tokens, None, min(n_frames, 8), language_id=-1, rep_penalty=1.0, temperature=0.0)
    np.savez_compressed(out, tokens=tokens, spk=spk, step_x=step_x,
                        prefill=pre.astype(np.float32), trailing=trail, pad=pad,
                        prefill_nospk=pre_ns, trailing_nospk=trail_ns,
                        prefill_hidden_last=hid[-1], prefill_logits=logits,
                        step_logits=step_logits, step_hidden=step_hidden,
                        cp_cb0=np.int32(cb0), cp_codes=cp_codes, codes=codes, codes_nospk=codes_nospk)
    ref.close()
    print("wrote", out, "frames", codes.shape)
    return codes


def gen_vocoder(codes, out, T):
    _, voc = models("tiny", "f16")
    dec = reflib.RefDecoder(voc)
    codes = np.ascontiguousarray(codes[:T] % 2048, np.int32)
    wav = dec.decode(codes)
    one = dec.decode(codes[:1])
    np.savez_compressed(out, codes=codes, wav=wav.astype(np.float32), wav_1frame=one.astype(np.float32))
    dec.close()
    print("wrote", out, wav.shape, "rms", float(np.sqrt((wav ** 2).mean())), "peak", float(np.abs(wav).max()))


def gen_legacy(out):
    """predict_codes (legacy parallel-heads graph, src/tts_transformer.cpp:2130-2180): logits of the 15
    heads for a seeded hidden vector, with and without previous codes, F16 and Q8_0 tiny models.  Only
    every 8th logit of each head is stored (bit patterns) to keep the fixture small."""
    rng = np.random.default_rng(77)
    hidden = (rng.standard_normal(1024) * 0.5).astype(np.float32)
    d = dict(hidden=hidden)
    for wtype in ("f16", "q8_0"):
        tts, _ = models("tiny", wtype)
        ref = reflib.RefTTS(tts)
        prev = rng.integers(0, ref.CPV, size=15).astype(np.int32)
        d[f"prev_{wtype}"] = prev
        d[f"logits_{wtype}"] = ref.predict_codes(hidden, prev)[:, ::8].copy()
        d[f"logits_noprev_{wtype}"] = ref.predict_codes(hidden, None)[:, ::8].copy()
        ref.close()
    np.savez_compressed(out, **d)
    print("wrote", out)


def gen_long(out, n_frames=4096):
    """BASELINE.json config 5 on the full-size F16 model: the reference's greedy codes for every one of the
    4096 frames (n_ctx 4114), same tokens / speaker vector as full_f16.npz.  ~51 ms per frame at 16 threads;
    stored as int16 (codes < 3072)."""
    tts, _ = models("full", "f16")
    ref = reflib.RefTTS(tts)
    spk, _ = golden_inputs()
    codes = ref.generate(HELLO_FULL, spk, n_frames, rep_penalty=1.05, temperature=0.0)
    ref.close()
    assert codes.max() < 32768 and codes.min() >= 0
    np.savez_compressed(out, tokens=HELLO_FULL, spk=spk, codes=codes.astype(np.int16))
    print("wrote", out, codes.shape)


def gen_vocoder_long(out, T=1024, n_win=32, win=2048):
    """Reference waveform for T = 1024 frames; only n_win windows of `win` samples are kept (the whole waveform is
    7.9 MB): enough to pin the multi-tile attention and the large-grid GEMM paths at a length the short goldens miss."""
    _, voc = models("tiny", "f16")
    dec = reflib.RefDecoder(voc)
    codes = np.random.default_rng(1024).integers(0, 2048, size=(T, 16), dtype=np.int32)
    wav = dec.decode(codes)
    dec.close()
    offsets = np.linspace(0, len(wav) - win, n_win).astype(np.int64)
    np.savez_compressed(out, codes=codes.astype(np.int16), offsets=offsets, win=np.stack([wav[o:o + win] for o in offsets]).astype(np.float32))
    print("wrote", out, wav.shape, "rms", float(np.sqrt((wav ** 2).mean())))


def resample_linear(x, in_rate, out_rate):
    """src/qwen3_tts.cpp:84-102 restated (double-precision position, float result)."""
    ratio = float(in_rate) / float(out_rate)
    n_out = int(float(len(x)) / ratio)
    src = np.arange(n_out, dtype=np.float64) * ratio
    i0 = src.astype(np.int64)
    frac = src - i0
    i1 = np.minimum(i0 + 1, len(x) - 1)
    y = ((1.0 - frac) * x[i0].astype(np.float64) + frac * x[i1].astype(np.float64)).astype(np.float32)
    y[i0 + 1 >= len(x)] = x[-1]
    return y


def gen_speaker(out, wav_path="/root/reference/examples/readme_clone_input.wav"):
    """Golden G7: the x-vector the reference's own ECAPA encoder (AudioTokenizerEncoder::encode) gives for
    examples/readme_clone_input.wav (16 kHz mono, 60 s, resampled to 24 kHz as the pipeline does) with the random-init
    spk_enc.* weights of tools/make_random_gguf.py.  Only the 1024 floats are stored, not the recording."""
    import wave
    with wave.open(wav_path, "rb") as w:
        assert w.getnchannels() == 1 and w.getsampwidth() == 2
        rate = w.getframerate()
        pcm = np.frombuffer(w.readframes(w.getnframes()), np.int16).astype(np.float32) / 32768.0
    x = resample_linear(pcm, rate, 24000)
    tts, _ = models("tiny", "f16")                       # both presets carry the same spk_enc.* tensors
    emb = reflib.speaker_embed(tts, x)
    assert emb.shape == (1024,) and np.isfinite(emb).all()
    np.savez_compressed(out, embedding=emb.astype(np.float32), n_samples_24k=np.int64(len(x)), source="examples/readme_clone_input.wav")
    print("wrote", out, "rms", float(np.sqrt((emb ** 2).mean())), "max", float(np.abs(emb).max()))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--full", action="store_true", help="also write the full-size (28-layer) goldens")
    ap.add_argument("--threads", type=int, default=8)
    ap.add_argument("--only-legacy", action="store_true", help="write only legacy_predict_codes.npz")
    ap.add_argument("--long", action="store_true", help="write only long_f16.npz (4096 greedy frames of the full F16 model)")
    ap.add_argument("--speaker", action="store_true", help="write only speaker_xvector.npz (golden G7)")
    ap.add_argument("--vocoder-long", action="store_true", help="write only vocoder_long_f16.npz (T = 1024 reference windows)")
    a = ap.parse_args()
    if not reflib.available():
        sys.exit("oracle/_ref/libq3t_ref.so missing: make -f oracle/Makefile.ref")
    reflib.RefLib.get().set_threads(a.threads)
    os.makedirs(GOLDEN, exist_ok=True)
    if a.speaker:
        gen_speaker(os.path.join(GOLDEN, "speaker_xvector.npz"))
        return
    if a.vocoder_long:
        gen_vocoder_long(os.path.join(GOLDEN, "vocoder_long_f16.npz"))
        return
    if a.long:
        gen_long(os.path.join(GOLDEN, "long_f16.npz"))
        return
    if a.only_legacy:
        gen_legacy(os.path.join(GOLDEN, "legacy_predict_codes.npz"))
        return
    codes = gen_tts("tiny", "f16", HELLO_TINY, 16, os.path.join(GOLDEN, "tiny_f16.npz"))
    gen_tts("tiny", "q8_0", HELLO_TINY, 16, os.path.join(GOLDEN, "tiny_q8_0.npz"))
    gen_vocoder(codes, os.path.join(GOLDEN, "vocoder_f16.npz"), 6)
    gen_legacy(os.path.join(GOLDEN, "legacy_predict_codes.npz"))
    if a.full:
        gen_tts("full", "f16", HELLO_FULL, 64, os.path.join(GOLDEN, "full_f16.npz"))
        gen_tts("full", "q8_0", HELLO_FULL, 64, os.path.join(GOLDEN, "full_q8_0.npz"))


if __name__ == "__main__":
    main()
