"""Streaming / batched front door (host/b200_pipeline.h) — B200 only.

The reference pipeline is one utterance at a time and decodes after the last frame (src/qwen3_tts.cpp:278-442);
the front door must return exactly what that order of operations returns: generate() -> decode().
"""
import numpy as np
import pytest

from util import HELLO_FULL, HELLO_TINY, models

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pair(q3t):
    t = q3t.TTSTransformer()
    d = q3t.AudioTokenizerDecoder()
    tts, voc = models("tiny", "f16")
    assert t.load_model(tts), t.get_error()
    assert d.load_model(voc), d.get_error()
    yield t, d
    t.close(); d.close()


@pytest.mark.parametrize("chunk", [1, 7, 32, 100])
def test_generate_in_chunks_equals_generate(pair, chunk):
    t, _ = pair
    rng = np.random.default_rng(chunk)
    spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
    want = t.generate(HELLO_TINY, spk, 45, temperature=0.0)
    parts = list(t.generate_stream(HELLO_TINY, spk, 45, chunk, temperature=0.0))
    assert all(len(p) <= chunk for p in parts)
    assert np.array_equal(np.concatenate(parts), want)
    # sampled: the same seed stream (set_seed restarts it) gives the same codes whatever the chunking
    t.set_seed(5)
    a = t.generate(HELLO_TINY, spk, 30, temperature=0.9, top_k=50)
    t.set_seed(5)
    b = np.concatenate(list(t.generate_stream(HELLO_TINY, spk, 30, chunk, temperature=0.9, top_k=50)))
    assert np.array_equal(a, b)


def test_synthesize_stream_equals_generate_then_decode_and_reports_first_audio(q3t, pair):
    t, d = pair
    rng = np.random.default_rng(1)
    spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
    codes = t.generate(HELLO_TINY, spk, 60, temperature=0.0)
    want = d.decode(codes)
    wav, st = q3t.synthesize_stream(t, d, HELLO_TINY, spk, 60, first_chunk=4, chunk=16)
    assert st.n_frames == 60 and st.n_chunks == 1 + (56 + 15) // 16 and st.n_samples == len(want) == len(wav)
    assert np.array_equal(wav, want)                          # streaming decode is bit-equal to the one-shot decode
    assert 0.0 < st.t_first_audio_ms < st.t_total_ms
    # the first audio arrives after 4 frames, not after 60
    wav2, st2 = q3t.synthesize_stream(t, d, HELLO_TINY, spk, 60, first_chunk=60, chunk=60)
    assert np.array_equal(wav2, want) and st2.n_chunks == 1
    assert st.t_first_audio_ms < st2.t_first_audio_ms


def test_synthesize_batch_equals_per_utterance_pipeline(q3t, pair):
    t, d = pair
    rng = np.random.default_rng(3)
    utts = []
    for i in range(11):                                       # one group of 8 and one of 3
        toks = HELLO_TINY.copy(); toks[3:11] = rng.integers(32, 120, size=8)
        utts.append((toks, None if i == 4 else (rng.standard_normal(1024) * 0.05).astype(np.float32)))
    wavs, st = q3t.synthesize_batch(t, d, utts, 24)
    assert len(wavs) == 11 and st.n_frames == 11 * 24 and st.n_samples == 11 * (1920 * 24 - 555)
    for (toks, spk), w in zip(utts, wavs):
        want = d.decode(t.generate(toks, spk, 24, temperature=0.0))
        assert np.array_equal(w, want)
    assert st.t_generate_ms > 0 and st.t_decode_ms > 0 and st.t_total_ms >= st.t_generate_ms
    # several decoder instances (one host thread and stream each): the same waveforms
    d2 = q3t.AudioTokenizerDecoder()
    assert d2.load_model(models("tiny", "f16")[1]), d2.get_error()
    wavs2, st2 = q3t.synthesize_batch(t, [d, d2], utts, 24)
    assert st2.n_samples == st.n_samples and all(np.array_equal(a, b) for a, b in zip(wavs, wavs2))
    wavs3 = q3t.decode_batch([d, d2], [t.generate(toks, spk, 24, temperature=0.0) for toks, spk in utts[:5]])
    assert all(np.array_equal(a, b) for a, b in zip(wavs[:5], wavs3))
    d2.close()
