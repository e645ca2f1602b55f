"""Developer script: a short program for `ncu --set full` captures — one batched generate (8 sequences,
lm_batch_kernel) and one vocoder decode (vgemm_umma) on the full-size random-init model."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models, HELLO_FULL
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 24
m = pkg()
tts, voc = models("full", "f16")
t = m.TTSTransformer()
assert t.load_model(tts), t.get_error()
rng = np.random.default_rng(5)
utts = [(HELLO_FULL, (rng.standard_normal(1024) * 0.05).astype(np.float32)) for _ in range(8)]
codes = t.generate_batch(utts, frames)
assert codes is not None, t.get_error()
d = m.AudioTokenizerDecoder()
assert d.load_model(voc), d.get_error()
big = np.concatenate([codes[0]] * max(1, 320 // frames))[:320]
wav = d.decode(big)
assert wav is not None, d.get_error()
print("ok", len(codes), codes[0].shape, wav.shape)
