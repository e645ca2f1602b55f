"""Developer script (not collected by pytest): per-phase clock breakdown of lm_kernel on a B200.
Arms the in-kernel profiler on one CTA, runs one frame program (16 code-predictor passes + talker
step) at a given context, and prints the mean cycles between consecutive marks."""
import os, sys, collections
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from util import pkg, models, HELLO_FULL

cta = int(sys.argv[1]) if len(sys.argv) > 1 else 0
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 40
m = pkg()
t = m.TTSTransformer()
assert t.load_model(models("full", "f16")[0]), t.get_error()
rng = np.random.default_rng(2024)
spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
t.stage_prompt(HELLO_FULL, spk)
t.run_staged(frames, with_vocoder=False)            # warm
t.stage_prompt(HELLO_FULL, spk)
t.profile_arm(cta, 1 << 18)
t.run_staged(frames, with_vocoder=False)
clk, tag = t.profile_read(1 << 18)
t.profile_arm(0, 0)
print("records", len(clk), "frames_ms", t.timing().frames_ms, "ms/frame", t.timing().frames_ms / frames)
# split into launches at clock resets is not needed: clock64 is monotonic per SM; a launch boundary shows as a big gap
pairs = collections.defaultdict(list)
for i in range(1, len(clk)):
    d = clk[i] - clk[i - 1]
    if d < 0 or d > 2_000_000:
        continue
    pairs[(int(tag[i - 1]), int(tag[i]))].append(d)
names = {1: "pass/layer start", 2: "rmsnormA done", 3: "waitA done", 4: "gemvA+store done", 5: "barA done", 6: "attention done",
         7: "barB done", 8: "loadC done", 9: "waitC done", 10: "gemvC done", 11: "barC done", 12: "load+rmsD done", 13: "waitD done",
         14: "gemvD done", 15: "barD done", 16: "loadE done", 17: "waitE done", 18: "gemvE done", 19: "barE done", 20: "layers done",
         21: "head done", 22: "bar head done", 30: "pass begin", 31: "x formed"}
tot = sum(sum(v) for v in pairs.values())
print(f"{'from':>18s} -> {'to':18s} {'count':>7s} {'mean cyc':>9s} {'share':>6s}")
for (a, b), v in sorted(pairs.items(), key=lambda kv: -sum(kv[1])):
    print(f"{names.get(a, a):>18} -> {names.get(b, b):18} {len(v):7d} {np.mean(v):9.0f} {sum(v) / tot * 100:5.1f}%")
