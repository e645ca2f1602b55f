"""Developer script (lives outside tests/): first-light comparison on a GPU box."""
import sys, time
import numpy as np
sys.path.insert(0, __import__('os').path.join(__import__('os').path.dirname(__import__('os').path.abspath(__file__)), '..', '..', 'tests'))
from util import pkg, models, HELLO_TINY, HELLO_FULL
from oracle import reflib

preset = sys.argv[1] if len(sys.argv) > 1 else "tiny"
nfr = int(sys.argv[2]) if len(sys.argv) > 2 else 12
tts_path, voc_path = models(preset)
toks = HELLO_TINY if preset == "tiny" else HELLO_FULL
m = pkg()
reflib.RefLib.get().set_threads(16)
ref = reflib.RefTTS(tts_path)
t = m.TTSTransformer()
assert t.load_model(tts_path), t.get_error()
rng = np.random.default_rng(0)
spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)

pre_r, trail_r, pad_r = ref.build_prompt(toks, spk)
pre, trail, pad = t.build_prompt(toks, spk)
print("prompt shapes", pre.shape, trail.shape, "max|d| prefill", np.abs(pre - pre_r).max(), "trailing", np.abs(trail - trail_r).max(), "pad", np.abs(pad - pad_r).max())

ref.init_kv(64); assert t.init_kv_cache(64), t.get_error()
hid_r, log_r = ref.prefill(pre_r)
out = t.forward_prefill(pre_r); assert out is not None, t.get_error()
hid, log = out
print("prefill hidden max|d|", np.abs(hid - hid_r).max(), "logits max|d|", np.abs(log - log_r).max(), "argmax", log.argmax(), log_r.argmax())

x = (rng.standard_normal(1024) * 0.05).astype(np.float32)
lr, hr = ref.step(x, pre_r.shape[0])
o = t.forward_step(x, pre_r.shape[0]); assert o is not None, t.get_error()
print("step logits max|d|", np.abs(o[0] - lr).max(), "hidden", np.abs(o[1] - hr).max())

cr = ref.predict_codes_ar(hr, int(lr[:2048].argmax()), 0.0, 0)
c = t.predict_codes_autoregressive(hr, int(lr[:2048].argmax()), 0.0, 0); assert c is not None, t.get_error()
print("codepred ref ", cr); print("codepred ours", c, "match", (c == cr).all())

t0 = time.time(); codes_r = ref.generate(toks, spk, nfr, temperature=0.0); tr = time.time() - t0
t0 = time.time(); codes = t.generate(toks, spk, nfr, temperature=0.0); tg = time.time() - t0
assert codes is not None, t.get_error()
same = codes.shape == codes_r.shape and (codes == codes_r).all()
print(f"generate {nfr} frames: ref {tr:.2f}s ours {tg:.3f}s  shape {codes.shape} identical={same}")
if not same:
    n = min(len(codes), len(codes_r)); bad = np.argwhere(codes[:n] != codes_r[:n]); print("first mismatch", bad[:3])
tm = t.timing(); print("timing prefill_ms", tm.prefill_ms, "frames_ms", tm.frames_ms, "ms/frame", tm.frames_ms / max(tm.n_frames, 1))
us = t.bench_talker_step(20, 64); wb = t.weight_bytes(1)
print(f"talker step {us:.1f} us, weights {wb/1e6:.1f} MB -> {wb/us/1e3:.1f} GB/s")
