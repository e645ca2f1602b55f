"""Sampled decoding (BASELINE.json config 3: temperature 0.9, top-k 50, repetition penalty 1.05) — B200 only.

The reference's RNG cannot be seeded (std::mt19937 from random_device, src/tts_transformer.h:336), so
sampled parity is what SURVEY.md §8c prescribes:
  (i)   the device sampler puts ZERO mass outside the reference's tie-inclusive top-k set,
  (ii)  its draw frequencies match the reference's probability vector (chi-square over 1e5 draws), for the
        code-predictor form and for the codebook-0 form with suppress + repetition penalty
        (oracle/port.py::sample_probs restates sample_or_argmax, src/tts_transformer.cpp:2349-2387, :2664-2722),
  (iii) teacher-forced along a trajectory the REFERENCE sampled on this box: every talker step's logits and
        hidden state are bit-equal, and every token the reference drew is inside our sampler's support.
"""
import os

import numpy as np
import pytest

from util import GOLDEN, HELLO_FULL, HELLO_TINY, models

pytestmark = pytest.mark.gpu

HAVE_REF = os.path.exists(os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "libq3t_ref.so"))


@pytest.fixture(scope="module")
def tts_tiny(q3t):
    t = q3t.TTSTransformer()
    assert t.load_model(models("tiny", "f16")[0]), t.get_error()
    yield t
    t.close()


def _chi2_ok(counts, probs, n):
    """Pearson chi-square against `probs`, bins with expectation < 5 pooled; accept below mean + 5 sigma."""
    exp = probs.astype(np.float64) * n
    big = exp >= 5.0
    o = np.concatenate([counts[big], [counts[~big].sum()]]).astype(np.float64)
    e = np.concatenate([exp[big], [exp[~big].sum()]])
    keep = e > 0
    o, e = o[keep], e[keep]
    chi2 = float(((o - e) ** 2 / e).sum())
    dof = len(e) - 1
    return chi2, dof, chi2 <= dof + 5.0 * np.sqrt(2.0 * dof) + 5.0


@pytest.mark.parametrize("vocab,top_k,temperature", [(2048, 50, 0.9), (2048, 7, 2.0), (2048, 0, 1.0), (3072, 50, 0.9)])
def test_sampler_distribution_vs_reference_probabilities(tts_tiny, vocab, top_k, temperature):
    from oracle import port
    rng = np.random.default_rng(vocab + top_k)
    logits = (rng.standard_normal(vocab) * 2.5).astype(np.float32)
    if top_k > 1:
        # ties exactly at the k-th largest value: the reference keeps all of them (`<` test)
        order = np.argsort(-logits)
        logits[order[top_k]] = logits[order[top_k - 1]]
        logits[order[top_k + 1]] = logits[order[top_k - 1]]
    process = vocab == 3072
    seen = rng.choice(2048, size=40, replace=False).astype(np.int32) if process else None
    if process:
        seen[:5] = np.argsort(-logits[:2048])[:5]          # penalise some of the likeliest tokens
        logits[2150] = logits[:2048].max() - 1.0           # EOS survives the suppression and stays drawable
    probs, lt = port.sample_probs(logits, temperature, top_k, process=process, seen=seen if process else (),
                                  rep_penalty=1.05 if process else 1.0)
    n = 100_000
    draws = tts_tiny.sample_logits(logits, n, temperature, top_k, seed=777, process=process, seen=seen,
                                   repetition_penalty=1.05 if process else 1.0)
    assert draws.min() >= 0 and draws.max() < vocab
    counts = np.bincount(draws, minlength=vocab)
    support = probs > 0
    assert counts[~support].sum() == 0, "mass outside the reference's threshold set"
    if top_k > 1 and not process:
        assert support.sum() == top_k + 2                  # the two ties at the threshold are kept
        assert counts[support].min() > 0 or probs[support].min() * n < 20
    if process:
        assert counts[2048:2150].sum() == 0 and counts[2151:].sum() == 0 and counts[2150] > 0
    chi2, dof, ok = _chi2_ok(counts, probs, n)
    assert ok, f"chi2 {chi2:.1f} for {dof} dof"
    # a different seed gives a different stream, the same seed the same stream
    again = tts_tiny.sample_logits(logits, 64, temperature, top_k, seed=777, process=process, seen=seen,
                                   repetition_penalty=1.05 if process else 1.0)
    assert np.array_equal(again, draws[:64])
    other = tts_tiny.sample_logits(logits, 64, temperature, top_k, seed=778, process=process, seen=seen,
                                   repetition_penalty=1.05 if process else 1.0)
    assert not np.array_equal(other, draws[:64])


def test_sampler_top_k_one_is_argmax(tts_tiny):
    rng = np.random.default_rng(1)
    logits = rng.standard_normal(2048).astype(np.float32)
    d = tts_tiny.sample_logits(logits, 256, 0.9, 1, seed=3)
    assert (d == int(logits.argmax())).all()


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not shipped")
@pytest.mark.parametrize("preset,wtype,tokens,frames", [("tiny", "f16", HELLO_TINY, 24), ("full", "q8_0", HELLO_FULL, 24)])
def test_teacher_forced_along_reference_sampled_trajectory(q3t, preset, wtype, tokens, frames):
    """The reference samples a trajectory (temp 0.9 / top-k 50 / rep 1.05, its own unseeded RNG); both sides are then
    driven along exactly those codes: talker step logits / hidden bit-equal at every frame, and each sampled token
    has non-zero probability under the device sampler's rules (cb0: suppress + repetition set; cb1: top-50 of head 0)."""
    from oracle import port, reflib
    tts_path = models(preset, wtype)[0]
    reflib.RefLib.get().set_threads(min(16, os.cpu_count() or 4))
    ref = reflib.RefTTS(tts_path)
    rng = np.random.default_rng(9)
    spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
    traj = ref.generate(tokens, spk, frames, rep_penalty=1.05, temperature=0.9, top_k=50)
    assert traj.shape[0] >= 4, "the reference stopped too early to test anything"
    pre, trail, pad = ref.build_prompt(tokens, spk)
    n_ctx = pre.shape[0] + traj.shape[0] + 8
    ref.init_kv(n_ctx)
    t = q3t.TTSTransformer()
    assert t.load_model(tts_path), t.get_error()
    assert t.init_kv_cache(n_ctx), t.get_error()
    p = port.PortTTS(tts_path)                       # embedding rows + code-predictor head 0 logits

    hr, lr = ref.prefill(pre)
    o = t.forward_prefill(pre)
    assert o is not None, t.get_error()
    assert np.array_equal(o[1], lr) and np.array_equal(o[0][-1], hr[-1])
    logits, hidden = o[1], o[0][-1]
    n_past = pre.shape[0]
    seen = []
    for f in range(traj.shape[0]):
        cb0 = int(traj[f, 0])
        probs, _ = port.sample_probs(logits, 0.9, 50, process=True, seen=seen, rep_penalty=1.05)
        assert probs[cb0] > 0, f"frame {f}: the reference drew cb0={cb0}, impossible under the restated rules"
        d = t.sample_logits(logits, 2000, 0.9, 50, seed=f, process=True, seen=np.array(seen, np.int32), repetition_penalty=1.05)
        assert (probs[d] > 0).all()                   # the device never leaves the reference's support
        seen.append(cb0)
        # code predictor, first code: head-0 logits given (hidden, cb0) from the bit-exact C oracle
        cb0_e = p._rows("talker.codec_embd.weight", [cb0], 1024)[0]
        hh = p._forward(p.cp, np.stack([hidden, cb0_e]), 0, f32_attn=1)
        l1 = p._linear("code_pred.lm_head.0.weight", hh[1], 2048, 1024)[0]
        p1, _ = port.sample_probs(l1, 0.9, 50)
        assert p1[int(traj[f, 1])] > 0
        # next talker input from the reference's codes (src/tts_transformer.cpp:2759-2790), then one step on both sides
        e = p._rows("talker.codec_embd.weight", [cb0], 1024)[0].copy()
        for cb in range(1, 16):
            e += p._rows(f"code_pred.codec_embd.{cb - 1}.weight", [int(traj[f, cb])], 1024)[0]
        e += trail[f] if f < trail.shape[0] else pad
        a = ref.step(e.astype(np.float32), n_past)
        b = t.forward_step(e.astype(np.float32), n_past)
        assert b is not None, t.get_error()
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), f"frame {f}: step not bit-exact"
        logits, hidden = b[0], b[1]
        n_past += 1
    ref.close()
    t.close()


def test_sampled_generate_runs_config3(q3t):
    """config 3 end to end on the tiny Q8_0 model: codes in range, seed-repeatable, seeds differ."""
    t = q3t.TTSTransformer()
    assert t.load_model(models("tiny", "q8_0")[0]), t.get_error()
    rng = np.random.default_rng(2)
    spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
    t.set_seed(42)
    a = t.generate(HELLO_TINY, spk, 48, repetition_penalty=1.05, temperature=0.9, top_k=50)
    t.set_seed(42)
    b = t.generate(HELLO_TINY, spk, 48, repetition_penalty=1.05, temperature=0.9, top_k=50)
    t.set_seed(43)
    c = t.generate(HELLO_TINY, spk, 48, repetition_penalty=1.05, temperature=0.9, top_k=50)
    assert a is not None and b is not None and c is not None, t.get_error()
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert a.min() >= 0 and a[:, 0].max() < 2048 and a[:, 1:].max() < 2048
    t.close()
