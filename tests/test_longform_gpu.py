"""Long contexts and long-form generation (BASELINE.json config 5) — B200 only.

Full-size model: tests/golden/long_f16.npz holds the reference's own greedy codes for ALL 4096 frames of config 5
(oracle/gen_golden.py --long, 28-layer F16 model, n_ctx 4114) and the 4096-frame run must reproduce every one of
them; teacher-forced steps behind 1030- and 4100-row prefills are compared bit-exact with the live reference for
the full-size F16 and Q8_0 models.  Tiny model: the same teacher-forced comparison at every attention slice
boundary (127..1030 rows).  The full-length vocoder decode is checked through size-independent properties
(1920*T - 555 finite samples, causal prefix equality).
"""
import os

import numpy as np
import pytest

from util import GOLDEN, models

pytestmark = pytest.mark.gpu

HAVE_REF = os.path.exists(os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "libq3t_ref.so"))


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not shipped")
@pytest.mark.parametrize("M", [127, 128, 129, 193, 300, 1030])
def test_long_context_steps_bit_exact_vs_live_reference(q3t, M):
    from oracle import reflib
    tts, _ = models("tiny", "f16")
    reflib.RefLib.get().set_threads(min(16, os.cpu_count() or 4))
    ref = reflib.RefTTS(tts)
    t = q3t.TTSTransformer()
    assert t.load_model(tts), t.get_error()
    rng = np.random.default_rng(M)
    x = (rng.standard_normal((M, 1024)) * 0.1).astype(np.float32)
    ref.init_kv(M + 8)
    assert t.init_kv_cache(M + 8), t.get_error()
    hr, lr = ref.prefill(x)
    o = t.forward_prefill(x)
    assert o is not None, t.get_error()
    assert np.array_equal(o[0], hr) and np.array_equal(o[1], lr)
    for s in range(4):
        xs = (rng.standard_normal(1024) * 0.1).astype(np.float32)
        a = ref.step(xs, M + s)
        b = t.forward_step(xs, M + s)
        assert b is not None, t.get_error()
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    ref.close()
    t.close()


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not shipped")
@pytest.mark.parametrize("wtype,M", [("f16", 1030), ("q8_0", 1030), ("f16", 4100)])
def test_full_size_long_context_steps_bit_exact_vs_live_reference(q3t, wtype, M):
    """The 28-layer model behind an M-row context: ring-fed K/V slices, the 128-row slice boundaries and the
    ATT_MAX_KV bookkeeping at n_kv in the thousands, against the reference run on this box."""
    from oracle import reflib
    tts, _ = models("full", wtype)
    reflib.RefLib.get().set_threads(min(32, os.cpu_count() or 4))
    ref = reflib.RefTTS(tts)
    t = q3t.TTSTransformer()
    assert t.load_model(tts), t.get_error()
    rng = np.random.default_rng(M)
    x = (rng.standard_normal((M, 1024)) * 0.1).astype(np.float32)
    ref.init_kv(M + 8)
    assert t.init_kv_cache(M + 8), t.get_error()
    hr, lr = ref.prefill(x)
    o = t.forward_prefill(x)
    assert o is not None, t.get_error()
    assert np.array_equal(o[1], lr), "prefill logits differ"
    assert np.array_equal(o[0][-1], hr[-1]) and np.array_equal(o[0][M // 2], hr[M // 2])
    for s in range(3):
        xs = (rng.standard_normal(1024) * 0.1).astype(np.float32)
        a = ref.step(xs, M + s)
        b = t.forward_step(xs, M + s)
        assert b is not None, t.get_error()
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), f"step {s} behind {M} rows differs"
    ref.close()
    t.close()


def test_long_form_4096_frames(q3t):
    g = np.load(os.path.join(GOLDEN, "full_f16.npz"))
    gl = np.load(os.path.join(GOLDEN, "long_f16.npz"))
    tts, voc = models("full", "f16")
    t = q3t.TTSTransformer()
    assert t.load_model(tts), t.get_error()
    a = t.generate(gl["tokens"], gl["spk"], 4096, temperature=0.0)
    assert a is not None, t.get_error()
    assert a.shape == (4096, 16)                                  # random weights never emit EOS
    assert np.array_equal(a[:64], g["codes"])
    want = gl["codes"].astype(np.int32)
    neq = np.argwhere(a != want)
    assert len(neq) == 0, f"{len(neq)} codes differ from the reference's 4096-frame run, first at frame/codebook {neq[0]}"
    assert a[:, 0].max() < 2048 and a[:, 1:].max() < 2048 and a.min() >= 0
    b = t.generate(g["tokens"], g["spk"], 4096, temperature=0.0)
    assert np.array_equal(a, b)
    t.close()
    d = q3t.AudioTokenizerDecoder()
    assert d.load_model(voc), d.get_error()
    wav = d.decode(a)
    assert wav is not None, d.get_error()
    assert wav.shape == (1920 * 4096 - 555,)
    assert np.isfinite(wav).all() and np.abs(wav).max() <= 1.0
    # causal vocoder: the first frames of the long decode equal a short decode of the same codes
    short = d.decode(a[:24])
    n = len(short) - 2400
    err = wav[:n].astype(np.float64) - short[:n].astype(np.float64)
    assert 10 * np.log10((short[:n].astype(np.float64) ** 2).sum() / max((err ** 2).sum(), 1e-300)) > 60.0
    d.close()
