"""Developer script for the wide attention path (lm_kernel.cu attention_wide).
  check [preset]: prefill M rows, then 8 single-token steps vs the live reference (bit-exact expected), M around the
                  Q3T_WIDE_MIN threshold and the 32/512-position boundaries of the P.V chains and V slices
  time:           talker step time of the full model vs n_kv (run once with Q3T_WIDE_MIN=0 and once without)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models
mode = sys.argv[1] if len(sys.argv) > 1 else "check"
m = pkg()
if mode == "time":
    tts, _ = models("full")
    t = m.TTSTransformer(); assert t.load_model(tts), t.get_error()
    assert t.init_kv_cache(4608)
    print("Q3T_WIDE_MIN", os.environ.get("Q3T_WIDE_MIN"))
    ns = [int(a) for a in sys.argv[2:]] or [40, 128, 200, 256, 257, 320, 400, 512, 768, 1024, 1536, 2048, 3072, 4096, 4600]
    for n in ns:
        print(f"n_kv {n + 1:5d}  talker step {t.bench_talker_step(n, 64):8.1f} us", flush=True)
    sys.exit(0)
from oracle import reflib
preset = sys.argv[2] if len(sys.argv) > 2 else "tiny"
tts, _ = models(preset)
reflib.RefLib.get().set_threads(16)
ref = reflib.RefTTS(tts)
t = m.TTSTransformer(); assert t.load_model(tts), t.get_error()
rng = np.random.default_rng(1)
bad = 0
for M in (100, 250, 256, 300, 505, 509, 1017, 1030, 2043):
    x = (rng.standard_normal((M, 1024)) * 0.1).astype(np.float32)
    ref.init_kv(M + 16); assert t.init_kv_cache(M + 16), t.get_error()
    hr, lr = ref.prefill(x); o = t.forward_prefill(x); assert o is not None, t.get_error()
    line = f"M={M:5d} prefill hid {np.abs(o[0]-hr).max():.3g} logits {np.abs(o[1]-lr).max():.3g} | steps:"
    for s in range(8):
        xs = (rng.standard_normal(1024) * 0.1).astype(np.float32)
        a = ref.step(xs, M + s); b = t.forward_step(xs, M + s); assert b is not None, t.get_error()
        d = max(float(np.abs(a[0] - b[0]).max()), float(np.abs(a[1] - b[1]).max()))
        bad += d != 0.0
        line += f" {d:.2g}"
    print(line, flush=True)
print("MISMATCHES", bad)
