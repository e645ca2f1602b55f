"""Developer script: a small run of every decode kernel for compute-sanitizer (tiny 2-layer model):
    compute-sanitizer --tool memcheck python tools/dev/dev_sanitize_target.py
batch-1 generate (greedy + sampled), a 13-utterance batched generate (two interleaved groups, batched prefill), vocoder."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models, HELLO_TINY

m = pkg()
tts_path, voc_path = models("tiny", "f16")
t = m.TTSTransformer()
assert t.load_model(tts_path), t.get_error()
rng = np.random.default_rng(0)
spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
c = t.generate(HELLO_TINY, spk, 6, temperature=0.0)
c2 = t.generate(HELLO_TINY, spk, 6, temperature=0.9, top_k=50)
utts = []
for i in range(13):
    toks = HELLO_TINY.copy(); toks[3:11] = rng.integers(32, 120, size=8)
    utts.append((toks, (rng.standard_normal(1024) * 0.05).astype(np.float32)))
b = t.generate_batch(utts, 5)
b2 = t.generate_batch(utts[:5], 4, temperature=0.9, top_k=50)
d = m.AudioTokenizerDecoder()
assert d.load_model(voc_path), d.get_error()
w = d.decode(c)
print("ok", c.shape, c2.shape, len(b), len(b2), w.shape)
