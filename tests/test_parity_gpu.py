"""Parity of the CUDA path (through the C ABI / host classes) with the reference — B200 only.

Checkers, in order of authority:
  * committed golden vectors written by the unmodified reference (tests/golden/, oracle/gen_golden.py);
  * the C restatement oracle on fresh seeded inputs (oracle/port.py);
  * the reference itself (oracle/_ref/libq3t_ref.so travels to the GPU box) on longer runs.
The decode kernels reproduce ggml-cpu's accumulation order, so logits / hidden states are compared
BIT-EXACT (the north-star tolerance is 1e-2 max-abs; exactness is what keeps greedy tokens
identical over hundreds of dependent steps) and codes must be identical.
"""
import os

import numpy as np
import pytest

from util import GOLDEN, HELLO_FULL, HELLO_TINY, models

pytestmark = pytest.mark.gpu

LOGIT_TOL = 1e-2      # BASELINE.json north_star: logits within 1e-2 max-abs
EOS_LIKE = 752        # the early-stop models' EOS row copies this token's row x 1.02 (tools/make_random_gguf.py)


def _gold(name):
    return np.load(os.path.join(GOLDEN, name))


@pytest.fixture(scope="module")
def tts_tiny(q3t):
    t = q3t.TTSTransformer()
    assert t.load_model(models("tiny", "f16")[0]), t.get_error()
    yield t
    t.close()


@pytest.fixture(scope="module")
def tts_full(q3t):
    t = q3t.TTSTransformer()
    assert t.load_model(models("full", "f16")[0]), t.get_error()
    yield t
    t.close()


@pytest.fixture(scope="module")
def tts_tiny_q8(q3t):
    t = q3t.TTSTransformer()
    assert t.load_model(models("tiny", "q8_0")[0]), t.get_error()
    yield t
    t.close()


@pytest.fixture(scope="module")
def tts_full_q8(q3t):
    t = q3t.TTSTransformer()
    assert t.load_model(models("full", "q8_0")[0]), t.get_error()
    yield t
    t.close()


def _fixture_for(request, preset, wtype):
    return request.getfixturevalue("tts_" + preset + ("_q8" if wtype == "q8_0" else ""))


def test_native_library_loaded(q3t):
    """The CUDA extension is the thing that runs: the in-tree .so is mapped into this process."""
    q3t.lib()
    maps = open("/proc/self/maps").read()
    assert "libq3t_b200.so" in maps


@pytest.mark.parametrize("wtype", ["f16", "q8_0"])
@pytest.mark.parametrize("preset", ["tiny", "full"])
def test_prompt_prefill_step_codepred_vs_golden(q3t, preset, wtype, request):
    t = _fixture_for(request, preset, wtype)
    g = _gold(f"{preset}_{wtype}.npz")
    pre, trail, pad = t.build_prompt(g["tokens"], g["spk"])
    assert np.array_equal(pre, g["prefill"]) and np.array_equal(trail, g["trailing"]) and np.array_equal(pad, g["pad"])
    pre_ns, trail_ns, _ = t.build_prompt(g["tokens"], None, -1)
    assert pre_ns.shape == (8, 1024)
    assert np.array_equal(pre_ns, g["prefill_nospk"]) and np.array_equal(trail_ns, g["trailing_nospk"])

    assert t.init_kv_cache(64), t.get_error()
    out = t.forward_prefill(g["prefill"])
    assert out is not None, t.get_error()
    hid, logits = out
    assert np.abs(logits - g["prefill_logits"]).max() <= LOGIT_TOL
    assert np.array_equal(logits, g["prefill_logits"]), "prefill logits not bit-exact"
    assert np.array_equal(hid[-1], g["prefill_hidden_last"])
    assert np.array_equal(t.get_hidden_states(), g["prefill_hidden_last"])

    out = t.forward_step(g["step_x"], 10)
    assert out is not None, t.get_error()
    assert np.abs(out[0] - g["step_logits"]).max() <= LOGIT_TOL
    assert np.array_equal(out[0], g["step_logits"]) and np.array_equal(out[1], g["step_hidden"])

    codes = t.predict_codes_autoregressive(g["step_hidden"], int(g["cp_cb0"]), 0.0, 0)
    assert codes is not None, t.get_error()
    assert np.array_equal(codes, g["cp_codes"])


@pytest.mark.parametrize("wtype", ["f16", "q8_0"])
@pytest.mark.parametrize("preset,n", [("tiny", 16), ("full", 64)])
def test_generate_greedy_codes_identical_vs_golden(q3t, preset, n, wtype, request):
    t = _fixture_for(request, preset, wtype)
    g = _gold(f"{preset}_{wtype}.npz")
    codes = t.generate(g["tokens"], g["spk"], n, repetition_penalty=1.05, temperature=0.0, top_k=0)
    assert codes is not None, t.get_error()
    assert codes.shape == g["codes"].shape
    assert np.array_equal(codes, g["codes"])
    codes = t.generate(g["tokens"], None, 8, language_id=-1, repetition_penalty=1.0, temperature=0.0)
    assert np.array_equal(codes, g["codes_nospk"])


def test_generate_is_prefix_stable_and_repeatable(tts_full):
    """Size-independent property at a longer length: a 400-frame greedy run starts with the golden 64
    frames (KV re-allocation / larger n_ctx must not change a bit) and repeats exactly."""
    g = _gold("full_f16.npz")
    a = tts_full.generate(g["tokens"], g["spk"], 400, temperature=0.0)
    assert a is not None and a.shape == (400, 16)
    assert np.array_equal(a[:64], g["codes"])
    b = tts_full.generate(g["tokens"], g["spk"], 400, temperature=0.0)
    assert np.array_equal(a, b)
    assert a[:, 0].max() < 2048 and not (a[:, 0] == 2150).any()          # suppressed ids / EOS never emitted
    assert a[:, 1:].max() < 2048 and a.min() >= 0


@pytest.mark.parametrize("wtype", ["f16", "q8_0"])
def test_step_vs_c_oracle_fresh_inputs(wtype, request):
    from oracle import port
    tts_tiny = _fixture_for(request, "tiny", wtype)
    p = port.PortTTS(models("tiny", wtype)[0])
    g = _gold(f"tiny_{wtype}.npz")
    rng = np.random.default_rng(11)
    p.init_kv(64)
    assert tts_tiny.init_kv_cache(64)
    p.prefill(g["prefill"])
    assert tts_tiny.forward_prefill(g["prefill"]) is not None
    n_past = 10
    for scale in (0.02, 0.3, 3.0):
        x = (rng.standard_normal(1024) * scale).astype(np.float32)
        a = p.step(x, n_past)
        b = tts_tiny.forward_step(x, n_past)
        assert b is not None, tts_tiny.get_error()
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
        cb0 = int(a[0][:2048].argmax())
        assert np.array_equal(p.predict_codes_ar(a[1], cb0), tts_tiny.predict_codes_autoregressive(a[1], cb0, 0.0, 0))
        n_past += 1


def test_prefill_lengths_vs_c_oracle(tts_tiny):
    """Ragged prefill lengths (1, 7, 8, 9, 33 rows) — the soft-max SIMD/tail split depends on the width."""
    from oracle import port
    p = port.PortTTS(models("tiny", "f16")[0])
    rng = np.random.default_rng(5)
    for m in (1, 7, 8, 9, 33):
        x = (rng.standard_normal((m, 1024)) * 0.1).astype(np.float32)
        p.init_kv(64)
        assert tts_tiny.init_kv_cache(64)
        a = p.prefill(x)
        b = tts_tiny.forward_prefill(x)
        assert b is not None, tts_tiny.get_error()
        assert np.array_equal(a[0], b[0]), f"hidden differs at M={m}"
        assert np.array_equal(a[1], b[1]), f"logits differ at M={m}"


def test_error_behaviour_matches_reference(q3t, tts_tiny):
    t = tts_tiny
    g = _gold("tiny_f16.npz")
    assert t.generate(g["tokens"][:3], None, 4) is None
    assert t.get_error() == "Need at least 4 text tokens for generation"
    out = t.generate(g["tokens"], g["spk"], 0)
    assert out is not None and out.shape == (0, 16)                  # max_len <= 0: success, empty
    assert t.init_kv_cache(16)
    assert t.forward_step(g["step_x"], 16) is None
    assert t.get_error() == "Context length exceeded"
    fresh = q3t.TTSTransformer()
    assert fresh.L.q3t_tts_generate(fresh.h, None, 0, None, 4, None, None, 2050, 1.05, 0.0, 0) == 0
    assert fresh.get_error() == "Model not loaded"
    assert fresh.load_model("/nonexistent.gguf") is False and fresh.get_error()
    fresh.close()


def test_sampler_semantics(tts_tiny):
    """temperature / top-k sampling (src/tts_transformer.cpp:2349-2387): top_k=1 must equal greedy,
    a fixed seed repeats, different seeds explore, samples stay inside the top-k set."""
    g = _gold("tiny_f16.npz")
    h, cb0 = g["step_hidden"], int(g["cp_cb0"])
    greedy = tts_tiny.predict_codes_autoregressive(h, cb0, 0.0, 0)
    t = tts_tiny
    t.set_seed(123)
    k1 = t.predict_codes_autoregressive(h, cb0, 0.9, 1)
    assert np.array_equal(k1, greedy)
    seen = set()
    for s in range(24):
        t.set_seed(1000 + s)
        c = t.predict_codes_autoregressive(h, cb0, 0.9, 50)
        assert c is not None and c.min() >= 0 and c.max() < 2048
        seen.add(int(c[0]))
    assert len(seen) > 3                                             # explores the top-50
    # first code: must be one of the 50 largest logits of lm_head[0] given (hidden, cb0)
    from oracle import port
    p = port.PortTTS(models("tiny", "f16")[0])
    cb0_e = p._rows("talker.codec_embd.weight", [cb0], 1024)[0]
    hh = p._forward(p.cp, np.stack([h, cb0_e]), 0, f32_attn=1)
    logits = p._linear("code_pred.lm_head.0.weight", hh[1], 2048, 1024)[0]
    top50 = set(np.argsort(-logits)[:50].tolist())
    assert seen <= top50
    # the k-th-largest threshold of the radix select, for small k: the first code stays inside the top-k set and
    # reaches more than one member of it
    for k in (2, 3, 7):
        topk = set(np.argsort(-logits)[:k].tolist())
        got = set()
        for s in range(40):
            t.set_seed(5000 + 97 * k + s)
            got.add(int(t.predict_codes_autoregressive(h, cb0, 2.0, k)[0]))
        assert got <= topk and len(got) > 1, (k, got, topk)


@pytest.mark.skipif(not os.path.exists(os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "libq3t_ref.so")),
                    reason="oracle/_ref not shipped")
def test_generate_320_frames_identical_to_live_reference(tts_full):
    """Config 1 of BASELINE.json: 'Hello, world!', greedy, 320 frames — every code identical to the
    reference's own ggml-CPU run on this box."""
    from oracle import reflib
    reflib.RefLib.get().set_threads(min(16, os.cpu_count() or 4))
    ref = reflib.RefTTS(models("full", "f16")[0])
    rng = np.random.default_rng(3)
    spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
    want = ref.generate(HELLO_FULL, spk, 320, temperature=0.0)
    ref.close()
    got = tts_full.generate(HELLO_FULL, spk, 320, temperature=0.0)
    assert got is not None and got.shape == want.shape
    neq = np.argwhere(got != want)
    assert len(neq) == 0, f"first mismatch at frame/codebook {neq[0]}"


@pytest.mark.skipif(not os.path.exists(os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "libq3t_ref.so")),
                    reason="oracle/_ref not shipped")
def test_generate_q8_0_320_frames_identical_to_live_reference(tts_full_q8):
    """Q8_0 layer weights (config 3's model): int8 x int8 block dots, activations re-quantised per
    mat-vec, ffn_down on de-quantised weights — greedy codes identical to the reference's ggml-CPU run."""
    from oracle import reflib
    reflib.RefLib.get().set_threads(min(16, os.cpu_count() or 4))
    ref = reflib.RefTTS(models("full", "q8_0")[0])
    rng = np.random.default_rng(4)
    spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
    want = ref.generate(HELLO_FULL, spk, 320, temperature=0.0)
    ref.close()
    got = tts_full_q8.generate(HELLO_FULL, spk, 320, temperature=0.0)
    assert got is not None and got.shape == want.shape
    neq = np.argwhere(got != want)
    assert len(neq) == 0, f"first mismatch at frame/codebook {neq[0]}"


@pytest.mark.parametrize("wtype", ["f16", "q8_0"])
def test_legacy_predict_codes_vs_golden_and_oracle(wtype, request):
    """predict_codes (src/tts_transformer.cpp:2130-2180, graph :1413-1532): 15 x 2048 logits,
    bit-exact against the reference's golden subset and against the C oracle on every logit, for the tiny
    and the full-size models; the call must leave the engine usable for generate()."""
    from oracle import port
    g = _gold("legacy_predict_codes.npz")
    t = _fixture_for(request, "tiny", wtype)
    a = t.predict_codes(g["hidden"], g[f"prev_{wtype}"])
    assert a is not None, t.get_error()
    assert a.shape == (15, 2048)
    assert np.array_equal(a[:, ::8], g[f"logits_{wtype}"])
    b = t.predict_codes(g["hidden"], None)
    assert np.array_equal(b[:, ::8], g[f"logits_noprev_{wtype}"])
    o = port.PortTTS(models("tiny", wtype)[0])
    assert np.array_equal(a, o.predict_codes(g["hidden"], g[f"prev_{wtype}"]))
    # full-size model, fresh inputs, against the oracle
    tf = _fixture_for(request, "full", wtype)
    of = port.PortTTS(models("full", wtype)[0])
    rng = np.random.default_rng(5)
    hid = (rng.standard_normal(1024) * 0.7).astype(np.float32)
    prev = rng.integers(0, 2048, size=15).astype(np.int32)
    got = tf.predict_codes(hid, prev)
    assert got is not None, tf.get_error()
    assert np.array_equal(got, of.predict_codes(hid, prev))
    bad = prev.copy(); bad[3] = 5000
    assert tf.predict_codes(hid, bad) is None and "out of range" in tf.get_error()
    gg = _gold(f"full_{wtype}.npz")
    codes = tf.generate(gg["tokens"], gg["spk"], 4, temperature=0.0)
    assert np.array_equal(codes, gg["codes"][:4])


HAVE_REF = os.path.exists(os.path.join(os.path.dirname(GOLDEN), "..", "oracle", "_ref", "libq3t_ref.so"))


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not shipped")
@pytest.mark.parametrize("wtype", ["f16", "q8_0"])
def test_legacy_wrappers_forward_text_and_forward_codec_vs_live_reference(wtype, request):
    """forward_text = project_text_tokens (+ speaker vector on every row) -> forward_prefill,
    forward_codec = codec_embd row -> forward_step (src/tts_transformer.cpp:1963-1991, 2110-2120); forward and
    forward_with_audio are forward_text without a speaker (:2853-2866).  Bit-exact against the reference."""
    from oracle import reflib
    t = _fixture_for(request, "tiny", wtype)
    reflib.RefLib.get().set_threads(min(16, os.cpu_count() or 4))
    ref = reflib.RefTTS(models("tiny", wtype)[0])
    g = _gold(f"tiny_{wtype}.npz")
    toks = g["tokens"][:12]
    for spk in (g["spk"], None):
        ref.init_kv(64)
        assert t.init_kv_cache(64)
        want = ref.forward_text(toks, spk)
        got = t.forward_text(toks, spk)
        assert got is not None, t.get_error()
        assert np.array_equal(got, want)
        for tok, n_past in ((7, 12), (2049, 13), (1500, 14)):
            a = ref.forward_codec(tok, n_past)
            b = t.forward_codec(tok, n_past)
            assert b is not None, t.get_error()
            assert np.array_equal(a, b), (tok, n_past)
    assert t.forward_codec(5000, 15) is None and "out of range" in t.get_error()
    ref.close()


@pytest.mark.skipif(not HAVE_REF, reason="oracle/_ref not shipped")
@pytest.mark.parametrize("wtype", ["f16", "q8_0"])
def test_eos_early_stop_identical_to_live_reference(q3t, wtype):
    """the EOS path.  tools/make_random_gguf.py eos_like makes the EOS row a louder copy of a token greedy
    decoding emits, so generation stops a few to a few dozen frames in; frame count and every code must equal the reference's
    (src/tts_transformer.cpp:2724-2729: EOS stops BEFORE a frame is emitted), for several utterances that stop at
    different frames, and the engine must be reusable afterwards (producer drain of in-flight bulk copies)."""
    from oracle import reflib
    tts = models("tiny", wtype, eos_like=EOS_LIKE)[0]
    reflib.RefLib.get().set_threads(min(16, os.cpu_count() or 4))
    ref = reflib.RefTTS(tts)
    t = q3t.TTSTransformer()
    assert t.load_model(tts), t.get_error()
    stops = []
    for i in range(8):
        rng = np.random.default_rng(100 + i)
        toks = HELLO_TINY.copy(); toks[3:11] = rng.integers(32, 120, size=8)
        spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
        want = ref.generate(toks, spk, 64, temperature=0.0)
        got = t.generate(toks, spk, 64, temperature=0.0)
        assert got is not None, t.get_error()
        assert got.shape == want.shape, f"utterance {i}: {got.shape[0]} frames, reference {want.shape[0]}"
        assert np.array_equal(got, want)
        stops.append(want.shape[0])
    assert any(0 < n < 64 for n in stops), f"no utterance stopped early: {stops} (re-tune EOS_LIKE)"
    assert len(set(stops)) > 1
    ref.close()
    t.close()


def test_two_devices_in_one_process(q3t):
    """kernel attributes (dynamic shared memory > 48 KB) are per device; a process that holds engines on
    two GPUs (QWEN3_TTS_DEVICE is read at every load_model) must be able to launch on both."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    g = _gold("tiny_f16.npz")
    tts, voc = models("tiny", "f16")
    outs, waves = [], []
    old = os.environ.get("QWEN3_TTS_DEVICE")
    try:
        objs = []
        for dev in (0, 1):
            os.environ["QWEN3_TTS_DEVICE"] = str(dev)
            t = q3t.TTSTransformer(); d = q3t.AudioTokenizerDecoder()
            assert t.load_model(tts), t.get_error()
            assert d.load_model(voc), d.get_error()
            objs.append((t, d))
        for t, d in objs:                                   # first launches on device 1 come after device 0's
            c = t.generate(g["tokens"], g["spk"], 16, temperature=0.0)
            assert c is not None, t.get_error()
            outs.append(c)
            w = d.decode(c)
            assert w is not None, d.get_error()
            waves.append(w)
        b = objs[1][0].generate_batch([(g["tokens"], g["spk"])] * 3, 8)
        assert b is not None and np.array_equal(b[0], g["codes"][:8])
        for t, d in objs:
            t.close(); d.close()
    finally:
        if old is None:
            os.environ.pop("QWEN3_TTS_DEVICE", None)
        else:
            os.environ["QWEN3_TTS_DEVICE"] = old
    assert np.array_equal(outs[0], g["codes"]) and np.array_equal(outs[1], g["codes"])
    assert np.array_equal(waves[0], waves[1])
