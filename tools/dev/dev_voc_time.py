"""Developer script: whole decodes of T frames, device time (events around the launches) and wall time, median of N."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models
T = int(sys.argv[1]) if len(sys.argv) > 1 else 320
N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
m = pkg(); d = m.AudioTokenizerDecoder(); assert d.load_model(models("full", "f16")[1]), d.get_error()
codes = np.random.default_rng(1).integers(0, 2048, size=(T, 16), dtype=np.int32)
for _ in range(3): d.decode(codes)
dev, wall = [], []
for _ in range(N):
    w0 = time.perf_counter(); wav = d.decode(codes); wall.append((time.perf_counter() - w0) * 1e3)
    t = m.binding.Timing(); d.L.q3t_get_timing(d.L.q3t_dec_engine(d.h), m.binding.C.byref(t)); dev.append(t.vocoder_ms)
print(f"T={T}: device median {np.median(dev):.3f} ms (min {min(dev):.3f}), wall median {np.median(wall):.3f} ms, checksum {float(np.abs(wav).sum()):.6f}")
