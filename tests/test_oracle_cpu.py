"""Pins the CPU restatement oracle (oracle/q3t_oracle.c driven by oracle/port.py).

 * against the golden vectors under tests/golden/, which oracle/gen_golden.py wrote by running
   the UNMODIFIED reference (ggml-CPU, AVX2 build) on seeded random-init GGUFs;
 * against the reference itself when oracle/_ref/libq3t_ref.so is present (this container).

The restatement follows the reference's accumulation order, so the bar is BIT-EXACT for logits
and hidden states, not a tolerance.
"""
import os

import numpy as np
import pytest

from util import GOLDEN, models

from oracle import port, reflib


def _gold(name):
    p = os.path.join(GOLDEN, name)
    if not os.path.exists(p):
        pytest.skip(f"{name} missing: python oracle/gen_golden.py")
    return np.load(p)


@pytest.fixture(scope="module")
def orc_tiny():
    tts, _ = models("tiny", "f16")
    return port.PortTTS(tts)


def test_scalar_helpers_known_answers():
    L = port.lib()
    # F16 round-to-nearest-even (ggml-impl.h): exact halves, ties, overflow to inf
    assert L.orc_r16(1.0) == 1.0
    assert L.orc_r16(1.00048828125) == 1.0            # tie -> even
    assert L.orc_r16(1.00146484375) == 1.001953125    # tie -> even (up)
    assert L.orc_r16(65520.0) == float("inf")
    assert L.orc_r16(5.960464477539063e-08) == 5.960464477539063e-08   # smallest subnormal
    import ctypes as C
    L.orc_v_expf.restype = C.c_float
    L.orc_v_expf.argtypes = [C.c_float]
    assert L.orc_v_expf(0.0) == 1.0
    assert abs(L.orc_v_expf(-1.0) - np.exp(-1.0)) < 1e-7
    assert L.orc_v_expf(float("-inf")) == 0.0


def test_prompt_assembly_matches_golden(orc_tiny):
    g = _gold("tiny_f16.npz")
    pre, trail, pad = orc_tiny.build_prompt(g["tokens"], g["spk"])
    assert pre.shape == g["prefill"].shape == (10, 1024)
    assert np.array_equal(pre, g["prefill"])
    assert np.array_equal(trail, g["trailing"])
    assert np.array_equal(pad, g["pad"])
    pre, trail, _ = orc_tiny.build_prompt(g["tokens"], None, -1)     # 8 rows: no speaker, no language
    assert pre.shape == (8, 1024)
    assert np.array_equal(pre, g["prefill_nospk"]) and np.array_equal(trail, g["trailing_nospk"])


def test_prefill_and_step_bit_exact_vs_golden(orc_tiny):
    g = _gold("tiny_f16.npz")
    orc_tiny.init_kv(64)
    hid, logits = orc_tiny.prefill(g["prefill"])
    assert np.array_equal(hid[-1], g["prefill_hidden_last"])
    assert np.array_equal(logits, g["prefill_logits"])
    sl, sh = orc_tiny.step(g["step_x"], 10)
    assert np.array_equal(sl, g["step_logits"]) and np.array_equal(sh, g["step_hidden"])
    codes = orc_tiny.predict_codes_ar(g["step_hidden"], int(g["cp_cb0"]))
    assert np.array_equal(codes, g["cp_codes"])


def test_generate_codes_identical_vs_golden(orc_tiny):
    g = _gold("tiny_f16.npz")
    codes = orc_tiny.generate(g["tokens"], g["spk"], 16, rep_penalty=1.05)
    assert codes.shape == (16, 16)
    assert np.array_equal(codes, g["codes"])
    codes = orc_tiny.generate(g["tokens"], None, 8, language_id=-1, rep_penalty=1.0)
    assert np.array_equal(codes, g["codes_nospk"])
    assert orc_tiny.generate(g["tokens"], g["spk"], 0).shape == (0, 16)       # max_len <= 0 -> empty


def test_q8_0_prefill_vs_golden():
    g = _gold("tiny_q8_0.npz")
    tts, _ = models("tiny", "q8_0")
    p = port.PortTTS(tts)
    p.init_kv(64)
    hid, logits = p.prefill(g["prefill"])
    assert np.array_equal(logits, g["prefill_logits"])
    codes = p.generate(g["tokens"], g["spk"], 6, rep_penalty=1.05)
    assert np.array_equal(codes, g["codes"][:6])


@pytest.mark.skipif(not reflib.available(), reason="oracle/_ref not built (make -f oracle/Makefile.ref)")
def test_live_reference_agrees_with_golden_and_oracle(orc_tiny):
    """The reference run here reproduces the committed goldens (they are machine independent) and
    the restatement matches it on a fresh random input."""
    g = _gold("tiny_f16.npz")
    tts, _ = models("tiny", "f16")
    reflib.RefLib.get().set_threads(4)
    ref = reflib.RefTTS(tts)
    ref.init_kv(64)
    hid, logits = ref.prefill(g["prefill"])
    assert np.array_equal(logits, g["prefill_logits"])
    rng = np.random.default_rng(7)
    x = (rng.standard_normal(1024) * 0.1).astype(np.float32)
    a = ref.step(x, 10)
    orc_tiny.init_kv(64)
    orc_tiny.prefill(g["prefill"])
    b = orc_tiny.step(x, 10)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    ref.close()


@pytest.mark.parametrize("wtype", ["f16", "q8_0"])
def test_legacy_predict_codes_matches_golden(wtype):
    """predict_codes (legacy parallel-heads graph, src/tts_transformer.cpp:2130-2180): every 8th logit of
    the 15 heads, with and without previous codes, bit-exact against the reference's output."""
    g = _gold("legacy_predict_codes.npz")
    t = port.PortTTS(models("tiny", wtype)[0])
    a = t.predict_codes(g["hidden"], g[f"prev_{wtype}"])
    assert a.shape == (15, 2048)
    assert np.array_equal(a[:, ::8], g[f"logits_{wtype}"])
    assert np.array_equal(t.predict_codes(g["hidden"], None)[:, ::8], g[f"logits_noprev_{wtype}"])
