"""Developer script (lives outside tests/): per-phase clock breakdown of lm_kernel on a B200.
Arms the in-kernel profiler (every CTA records its own marks with plain stores), runs a few frames
and prints, for the LAST frame launch, the mean cycles between consecutive marks for some CTAs and
the min / max over CTAs of every segment (a CTA that never waits in a poll is the critical one)."""
import os, sys, collections
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models, HELLO_FULL

frames = int(sys.argv[1]) if len(sys.argv) > 1 else 40
print("POLL_D0", os.environ.get("Q3T_POLL_D0"))
m = pkg()
t = m.TTSTransformer()
assert t.load_model(models("full", os.environ.get("Q3T_WTYPE", "f16"))[0]), t.get_error()
rng = np.random.default_rng(2024)
spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
nbatch = int(sys.argv[2]) if len(sys.argv) > 2 else 0
if nbatch:
    utts = [(HELLO_FULL, (rng.standard_normal(1024) * 0.05).astype(np.float32)) for _ in range(nbatch)]
    t.generate_batch(utts, frames)
    t.profile_arm(-1, 4096)
    t.generate_batch(utts, frames)
else:
    t.stage_prompt(HELLO_FULL, spk)
    t.run_staged(frames, with_vocoder=False)            # warm
    t.stage_prompt(HELLO_FULL, spk)
    t.profile_arm(frames // 2, 4096)
    t.run_staged(frames, with_vocoder=False)
print("frames_ms", t.timing().frames_ms, "ms/frame", t.timing().frames_ms / frames)
names = {1: "layer start", 2: "x polled", 3: "rmsnormA done", 4: "sliceA ready", 5: "gemvA+store", 6: "attention done",
         7: "att polled", 8: "sliceO ready", 9: "gemvC+store", 10: "x1 polled", 11: "rmsnormD done", 12: "sliceGU ready",
         13: "gemvD+store", 14: "m polled", 15: "sliceD ready", 16: "gemvE+store", 20: "head norm done", 21: "head done",
         **{100 + k: v + " g1" for k, v in {3: "rmsnormA done", 5: "gemvA+store", 7: "att polled", 9: "gemvC+store", 11: "rmsnormD done", 13: "gemvD+store", 15: "E done", 16: "gemvE+store"}.items()},
         50: "att stage0 done", 51: "att qkv in (w0)", 52: "att prologue bar", 53: "att scores done", 54: "att P ready", 55: "att V slices done", 56: "att stored", 57: "att lo begin", 58: "att lo end", 59: "att K slice in", 40: "gemvA warp0", 41: "gemvA cbar", 42: "qkv stored", 30: "pass begin", 31: "x formed", 60: "E poll0", 67: "E acquire", 61: "E comp0", 62: "E poll1", 63: "E comp1", 64: "E poll2", 65: "E comp2", 60: "A acquired", 61: "D acquired", **{80 + k: f"x in ({k} polls)" for k in range(10)}, **{90 + k: f"x1 in ({k} polls)" for k in range(10)}, 70: "norm polled", 71: "norm scale", 74: "att qkv poll begin", 75: "att qkv polled", 72: "att K scored (w0)", 73: "att K released"}
per_cta = {}
for cta in range(148):
    clk, tag = t.profile_read(4096, cta)
    pairs = collections.defaultdict(list)
    for i in range(1, len(clk)):
        pairs[(int(tag[i - 1]), int(tag[i]))].append(int(clk[i] - clk[i - 1]))
    per_cta[cta] = (pairs, int(clk[-1] - clk[0]) if len(clk) else 0, len(clk))
t.profile_arm(0, 0)
print("records per CTA", per_cta[0][2], "kernel cycles CTA0", per_cta[0][1], "CTA77", per_cta[77][1])
keys = sorted(per_cta[0][0].keys(), key=lambda k: -sum(per_cta[0][0][k]))
tot0 = sum(sum(v) for v in per_cta[0][0].values())
print(f"{'from':>16s} -> {'to':16s} {'count':>6s} {'CTA0':>7s} {'CTA20':>7s} {'CTA77':>7s} {'CTA147':>7s} {'min/CTA':>8s} {'max/CTA':>8s} {'share0':>6s}")
for k in keys:
    def mean(c):
        v = per_cta[c][0].get(k, [0]); return np.mean(v)
    allm = [mean(c) for c in range(148)]
    print(f"{names.get(k[0], k[0]):>16} -> {names.get(k[1], k[1]):16} {len(per_cta[0][0][k]):6d} {mean(0):7.0f} {mean(20):7.0f} {mean(77):7.0f} {mean(147):7.0f} "
          f"{min(allm):8.0f} {max(allm):8.0f} {sum(per_cta[0][0][k]) / tot0 * 100:5.1f}%")
