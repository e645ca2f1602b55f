"""Batched generation (config 4's inner loop: several utterances advancing in lock-step on one GPU).

The batched kernel keeps, per sequence, exactly the accumulation order of the batch-1 kernel, which
is itself bit-exact with the reference (tests/test_parity_gpu.py): the codes of every utterance must
equal what generate() returns for it alone, for any batch size and any mix of prompt lengths.
"""
import numpy as np
import pytest

from util import HELLO_FULL, models

pytestmark = pytest.mark.gpu


def _utterances(n, seed=17, speakers="mixed"):
    """speakers: "mixed" = every fifth utterance without a speaker vector (9 prompt rows instead of 10: such a wave is
    prefilled one sequence at a time), "all" / "none" = uniform waves (prefilled in lock-step by the batched kernel)."""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        body = rng.integers(32, 127, size=int(rng.integers(8, 65)), dtype=np.int32)       # 8..64 byte tokens
        toks = np.concatenate([HELLO_FULL[:3], body, HELLO_FULL[-5:]]).astype(np.int32)
        spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
        if speakers == "none" or (speakers == "mixed" and i % 5 == 4):
            spk = None
        out.append((toks, spk))
    return out


@pytest.fixture(scope="module")
def tts(q3t):
    t = q3t.TTSTransformer()
    assert t.load_model(models("full", "f16")[0]), t.get_error()
    yield t
    t.close()


# up to 8 sequences run as one lock-step group, 9..16 as two interleaved groups (11 = 6 + 5, 16 = 8 + 8), more in waves of 16
@pytest.mark.parametrize("n,frames", [(2, 24), (8, 40), (11, 24), (16, 40), (21, 16)])
def test_batched_codes_equal_single_sequence_codes(tts, n, frames):
    utts = _utterances(n, seed=100 + n)
    got = tts.generate_batch(utts, frames)
    assert got is not None, tts.get_error()
    assert len(got) == n
    for i, (toks, spk) in enumerate(utts):
        want = tts.generate(toks, spk, frames, temperature=0.0)
        assert want is not None, tts.get_error()
        assert got[i].shape == want.shape, f"utterance {i}"
        neq = np.argwhere(got[i] != want)
        assert len(neq) == 0, f"utterance {i}: first mismatch at frame/codebook {neq[0]}"


@pytest.mark.parametrize("n,speakers", [(13, "all"), (5, "none"), (19, "all")])
def test_batched_prefill_waves_equal_single_sequence_codes(tts, n, speakers):
    """Waves whose sequences all have the same number of prompt rows are prefilled in lock-step by the batched kernel
    (prompt rows as XM_BUF passes with the prefill's soft-max row width) instead of one sequence at a time on the batch-1
    kernel: the codes still equal generate() for each utterance alone."""
    utts = _utterances(n, seed=500 + n, speakers=speakers)
    got = tts.generate_batch(utts, 16)
    assert got is not None, tts.get_error()
    for i, (toks, spk) in enumerate(utts):
        want = tts.generate(toks, spk, 16, temperature=0.0)
        assert want is not None, tts.get_error()
        assert got[i].shape == want.shape, f"utterance {i}"
        neq = np.argwhere(got[i] != want)
        assert len(neq) == 0, f"utterance {i}: first mismatch at frame/codebook {neq[0]}"


def test_batched_q8_0_codes_equal_single_sequence_codes(q3t):
    """Q8_0 layer weights in the batched engine (int8 block dots, activations re-quantised per mat-vec exactly like the
    batch-1 kernel): 13 utterances = two interleaved groups (7 + 6); codes equal generate() for each alone, which
    tests/test_parity_gpu.py pins to the live reference for 320 frames."""
    t = q3t.TTSTransformer()
    assert t.load_model(models("full", "q8_0")[0]), t.get_error()
    utts = _utterances(13, seed=813)
    got = t.generate_batch(utts, 24)
    assert got is not None, t.get_error()
    for i, (toks, spk) in enumerate(utts):
        want = t.generate(toks, spk, 24, temperature=0.0)
        assert want is not None, t.get_error()
        assert got[i].shape == want.shape, f"utterance {i}"
        neq = np.argwhere(got[i] != want)
        assert len(neq) == 0, f"utterance {i}: first mismatch at frame/codebook {neq[0]}"
    t.close()


def test_batched_sampled_codes_equal_single_sequence_sampled_codes(tts):
    """Sampled decoding inside the batched engine (temperature 0.9, top-k 50, repetition penalty 1.05 = config 3's
    settings): utterance i draws from the stream seeded seed + i with the same routine as the batch-1 kernel, so its
    codes equal the engine's single-sequence run with that seed (whose sampler is pinned against the reference's
    distribution and its teacher-forced trajectory in tests/test_sampling_gpu.py)."""
    utts = _utterances(11, seed=311)
    tts.set_seed(1234)
    got = tts.generate_batch(utts, 24, temperature=0.9, top_k=50)
    assert got is not None, tts.get_error()
    greedy = tts.generate_batch(utts, 24)
    n_diff = 0
    for i, (toks, spk) in enumerate(utts):
        tts.stage_prompt(toks, spk)
        nf, _ = tts.run_staged(24, repetition_penalty=1.05, temperature=0.9, top_k=50, seed=1234 + i, with_vocoder=False)
        want = tts.fetch_codes(nf)
        assert got[i].shape == want.shape, f"utterance {i}"
        neq = np.argwhere(got[i] != want)
        assert len(neq) == 0, f"utterance {i}: first mismatch at frame/codebook {neq[0]}"
        n_diff += int(not np.array_equal(got[i], greedy[i]))
    assert n_diff >= 10                                          # it really sampled


def test_config4_job_64_utterances_320_frames_equals_generate(tts):
    """The job bench.py times for config 4 (64 utterances x 320 frames through the batched engine): every code of
    every utterance equals generate() for it alone -- and generate() at 320 frames is compared with the live
    reference in tests/test_parity_gpu.py."""
    utts = _utterances(64, seed=4)
    got = tts.generate_batch(utts, 320)
    assert got is not None, tts.get_error()
    assert len(got) == 64
    for i, (toks, spk) in enumerate(utts):
        want = tts.generate(toks, spk, 320, temperature=0.0)
        assert want is not None, tts.get_error()
        assert got[i].shape == want.shape == (320, 16), f"utterance {i}"
        neq = np.argwhere(got[i] != want)
        assert len(neq) == 0, f"utterance {i}: first mismatch at frame/codebook {neq[0]}"


@pytest.mark.parametrize("n_utt", [11, 19])
def test_batched_early_stop_two_waves(q3t, n_utt):
    """EOS inside the batched engine: utterances of the EOS-scaled model, 11 = one launch of two interleaved
    groups (6 + 5), 19 = a wave of 16 (8 + 8) + a wave of 3 re-using the buffers; sequences stop at different frames
    (finished ones keep flowing with frozen outputs), frame counts and codes equal generate() alone (itself pinned to
    the reference in test_parity_gpu)."""
    from test_parity_gpu import EOS_LIKE
    from util import HELLO_TINY
    t = q3t.TTSTransformer()
    assert t.load_model(models("tiny", "f16", eos_like=EOS_LIKE)[0]), t.get_error()
    utts = []
    for i in range(n_utt):
        rng = np.random.default_rng(100 + i)
        toks = HELLO_TINY.copy(); toks[3:11] = rng.integers(32, 120, size=8)
        utts.append((toks, (rng.standard_normal(1024) * 0.05).astype(np.float32)))
    got = t.generate_batch(utts, 64)
    assert got is not None, t.get_error()
    stops = []
    for i, (toks, spk) in enumerate(utts):
        want = t.generate(toks, spk, 64, temperature=0.0)
        assert got[i].shape == want.shape, f"utterance {i}: {got[i].shape[0]} frames vs {want.shape[0]}"
        assert np.array_equal(got[i], want), f"utterance {i}"
        stops.append(want.shape[0])
    assert len(set(stops)) > 2 and any(n < 64 for n in stops), stops
    again = t.generate_batch(utts[:3], 64)                      # the engine is reusable after early stops
    assert all(np.array_equal(again[i], got[i]) for i in range(3))
    t.close()


def test_batch_errors(tts):
    assert tts.generate_batch([(HELLO_FULL[:3], None)], 8) is None
    assert tts.get_error() == "Need at least 4 text tokens for generation"
    assert tts.generate_batch([(HELLO_FULL, None)], 0) == [] or all(len(c) == 0 for c in tts.generate_batch([(HELLO_FULL, None)], 0))
