"""Developer script: vocoder alone on random codes (for ncu launch lists)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models
T = int(sys.argv[1]) if len(sys.argv) > 1 else 320
m = pkg(); d = m.AudioTokenizerDecoder(); assert d.load_model(models("full", "f16")[1]), d.get_error()
codes = np.random.default_rng(1).integers(0, 2048, size=(T, 16), dtype=np.int32)
for i in range(3):
    w0 = time.perf_counter(); wav = d.decode(codes); print("decode ms", (time.perf_counter() - w0) * 1e3, len(wav), flush=True)
