"""Developer script: config 4 on one GPU (64 utterances through generate_batch) with the call's wall-clock
split (Q3T_BATCH_TRACE=1) next to the CUDA-event times."""
import os, sys, time, importlib
os.environ.setdefault("Q3T_BATCH_TRACE", "1")
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 320
m = pkg(); t = m.TTSTransformer()
assert t.load_model(models("full", "f16")[0]), t.get_error()
disp = importlib.import_module("q3t_b200.dispatch")
utts = disp.make_workload(64, seed=4, max_len=frames)
pairs = [(u.tokens, u.spk) for u in utts]
assert t.generate_batch(pairs[:8], 8) is not None
for rep in range(2):
    w0 = time.perf_counter(); codes = t.generate_batch(pairs, frames, repetition_penalty=1.05); w = time.perf_counter() - w0
    tm = t.timing(); nf = sum(len(c) for c in codes)
    print(f"rep {rep}: wall {w:.3f} s, events prefill {tm.prefill_ms:.1f} ms frames {tm.frames_ms:.1f} ms, {nf * 0.08 / w:.1f} audio-s/s codes only", flush=True)
