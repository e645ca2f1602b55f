"""Developer script: throughput of generate_batch (frames_ms from CUDA events) for several batch sizes."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models, HELLO_FULL
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 120
sizes = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 2, 4, 8, 16]
wave = int(os.environ.get("Q3T_BATCH_WAVE", "16"))
m = pkg(); t = m.TTSTransformer()
assert t.load_model(models("full", "f16")[0]), t.get_error()
rng = np.random.default_rng(7)
def utts(n):
    out = []
    for i in range(n):
        body = rng.integers(32, 127, size=int(rng.integers(8, 65)), dtype=np.int32)
        out.append((np.concatenate([HELLO_FULL[:3], body, HELLO_FULL[-5:]]).astype(np.int32), (rng.standard_normal(1024) * 0.05).astype(np.float32)))
    return out
for n in sizes:
    u = utts(n)
    t.generate_batch(u, 8)
    w0 = time.perf_counter(); c = t.generate_batch(u, frames); w = time.perf_counter() - w0
    tm = t.timing()
    nf = sum(len(x) for x in c)
    print(f"batch {n:3d}: wall {w*1e3:8.1f} ms  prefill {tm.prefill_ms:7.1f} frames {tm.frames_ms:8.1f} ms  -> {tm.frames_ms / frames / ((n + wave - 1) // wave):6.3f} ms per batched frame (waves of {wave}), "
          f"{nf * 0.08 / w:7.1f} audio-s/s wall, {nf * 0.08 / (tm.frames_ms * 1e-3):7.1f} audio-s/s frames only", flush=True)
