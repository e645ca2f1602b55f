"""Developer script: 64 utterances x T frames decoded by 1..4 decoder instances (q3t_pipe_decode_batch_multi), wall time per job."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models
T = int(sys.argv[1]) if len(sys.argv) > 1 else 320
N = int(sys.argv[2]) if len(sys.argv) > 2 else 64
m = pkg()
rng = np.random.default_rng(1)
codes = [rng.integers(0, 2048, size=(T, 16), dtype=np.int32) for _ in range(N)]
decs = []
for k in range(4):
    d = m.AudioTokenizerDecoder(); assert d.load_model(models("full", "f16")[1]), d.get_error()
    d.decode(codes[0]); decs.append(d)
for rep in range(2):
    for nd in (1, 2, 3, 4):
        w0 = time.perf_counter(); w = m.decode_batch(decs[:nd], codes); dt = time.perf_counter() - w0
        print(f"decoders {nd}: {dt*1e3:.1f} ms for {N} x {T} frames ({dt*1e3/N:.2f} ms per utterance)", flush=True)
w0 = time.perf_counter()
for c in codes: decs[0].decode(c)
print(f"python loop, 1 decoder: {(time.perf_counter()-w0)*1e3:.1f} ms", flush=True)
