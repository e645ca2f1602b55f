"""Developer script: wall clock vs CUDA-event time of run_staged / generate (host overhead hunt)."""
import os, sys, time, subprocess, threading
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models, HELLO_FULL
frames = int(sys.argv[1]) if len(sys.argv) > 1 else 320
poll = len(sys.argv) > 2 and sys.argv[2] == "smi"
m = pkg(); t = m.TTSTransformer()
assert t.load_model(models("full", "f16")[0]), t.get_error()
d = m.AudioTokenizerDecoder(); assert d.load_model(models("full", "f16")[1])
rng = np.random.default_rng(2024)
spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
stop = threading.Event()
def loop():
    while not stop.is_set():
        subprocess.run(["nvidia-smi", "--id=0", "--query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown", "--format=csv,noheader,nounits"], capture_output=True)
        stop.wait(0.2)
if poll:
    threading.Thread(target=loop, daemon=True).start()
t.stage_prompt(HELLO_FULL, spk)
for i in range(6):
    w0 = time.perf_counter(); t.run_staged(frames, with_vocoder=True); w = time.perf_counter() - w0
    tm = t.timing()
    print(f"staged {i}: wall {w*1e3:8.1f} ms  events prefill {tm.prefill_ms:6.1f} frames {tm.frames_ms:7.1f} voc {tm.vocoder_ms:6.1f}", flush=True)
for i in range(4):
    w0 = time.perf_counter(); c = t.generate(HELLO_FULL, spk, frames, temperature=0.0); w1 = time.perf_counter(); wav = d.decode(c); w2 = time.perf_counter()
    print(f"e2e {i}: generate {1e3*(w1-w0):8.1f} ms decode {1e3*(w2-w1):7.1f} ms", flush=True)
stop.set()
