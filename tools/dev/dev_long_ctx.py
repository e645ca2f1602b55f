"""Developer script: talker steps at long contexts vs the live reference (bit-exact check)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models
from oracle import reflib
preset = sys.argv[1] if len(sys.argv) > 1 else "tiny"
tts, _ = models(preset)
reflib.RefLib.get().set_threads(16)
ref = reflib.RefTTS(tts)
m = pkg(); t = m.TTSTransformer(); assert t.load_model(tts), t.get_error()
rng = np.random.default_rng(1)
for M in (20, 31, 32, 40, 49, 100, 193, 300, 1030):
    x = (rng.standard_normal((M, 1024)) * 0.1).astype(np.float32)
    ref.init_kv(M + 16); assert t.init_kv_cache(M + 16), t.get_error()
    hr, lr = ref.prefill(x); o = t.forward_prefill(x); assert o is not None, t.get_error()
    line = f"M={M:5d} prefill hid {np.abs(o[0]-hr).max():.3g} logits {np.abs(o[1]-lr).max():.3g} | steps:"
    for s in range(8):
        xs = (rng.standard_normal(1024) * 0.1).astype(np.float32)
        a = ref.step(xs, M + s); b = t.forward_step(xs, M + s); assert b is not None, t.get_error()
        line += f" {np.abs(a[0]-b[0]).max():.2g}"
    print(line, flush=True)
