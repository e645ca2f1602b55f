"""Batched generation (config 4's inner loop: several utterances advancing in lock-step on one GPU).

The batched kernel keeps, per sequence, exactly the accumulation order of the batch-1 kernel, which
is itself bit-exact with the reference (tests/test_parity_gpu.py): the codes of every utterance must
equal what generate() returns for it alone, for any batch size and any mix of prompt lengths.
"""
import numpy as np
import pytest

from util import HELLO_FULL, models

pytestmark = pytest.mark.gpu


def _utterances(n, seed=17):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        body = rng.integers(32, 127, size=int(rng.integers(8, 65)), dtype=np.int32)       # 8..64 byte tokens
        toks = np.concatenate([HELLO_FULL[:3], body, HELLO_FULL[-5:]]).astype(np.int32)
        spk = None if i % 5 == 4 else (rng.standard_normal(1024) * 0.05).astype(np.float32)
        out.append((toks, spk))
    return out


@pytest.fixture(scope="module")
def tts(q3t):
    t = q3t.TTSTransformer()
    assert t.load_model(models("full", "f16")[0]), t.get_error()
    yield t
    t.close()


@pytest.mark.parametrize("n,frames", [(2, 24), (8, 40), (11, 24)])
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


def test_batch_errors(tts):
    assert tts.generate_batch([(HELLO_FULL[:3], None)], 8) is None
    assert tts.get_error() == "Need at least 4 text tokens for generation"
    assert tts.generate_batch([(HELLO_FULL, None)], 0) == [] or all(len(c) == 0 for c in tts.generate_batch([(HELLO_FULL, None)], 0))
