"""Developer script: ms/frame of a short staged run (for parameter sweeps via environment variables)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models, HELLO_FULL
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 60
m = pkg(); t = m.TTSTransformer()
assert t.load_model(models("full", os.environ.get("Q3T_WTYPE", "f16"))[0]), t.get_error()
rng = np.random.default_rng(2024)
spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
for _ in range(3):
    t.stage_prompt(HELLO_FULL, spk); t.run_staged(frames, with_vocoder=False)
print({k: v for k, v in os.environ.items() if k.startswith("Q3T_")}, "ms/frame", t.timing().frames_ms / frames, "talker_us", t.bench_talker_step(40, 32))
