"""Developer script (lives outside tests/): per-stage comparison of the CUDA vocoder against the
reference's ggml graph nodes.  Reference side runs anywhere; the CUDA side needs a B200."""
import os, sys, tempfile
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
from util import pkg, models
from oracle import reflib

T = int(sys.argv[1]) if len(sys.argv) > 1 else 2
_, voc = models("tiny")
rng = np.random.default_rng(100 + T)
codes = rng.integers(0, 2048, size=(T, 16), dtype=np.int32)
reflib.RefLib.get().set_threads(8)
ref = reflib.RefDecoder(voc)
tr = reflib.trace_decoder(ref, codes)
named = {}
for k, (buf, ne) in tr.items():
    idx, op, nm = k.split(":", 2)
    named.setdefault(nm, []).append((int(idx), op, buf, ne))
if len(sys.argv) > 2 and sys.argv[2] == "list":
    for k, (buf, ne) in tr.items():
        print(k, ne, float(np.abs(buf).max()))
    sys.exit(0)

def ref_node(name):            # ggml [ne0 fastest] -> numpy
    idx, op, buf, ne = named[name][-1]
    return buf.reshape([n for n in reversed(ne)]).squeeze()

m = pkg()
d = m.AudioTokenizerDecoder()
dbg = tempfile.mkdtemp()
os.environ["Q3T_VOC_DEBUG_DIR"] = dbg
assert d.load_model(voc), d.get_error()
wav = d.decode(codes)
want = ref.decode(codes)
def snr(a, b):
    a = np.asarray(a, np.float64).ravel(); b = np.asarray(b, np.float64).ravel()
    return 10 * np.log10((a ** 2).sum() / max(((a - b) ** 2).sum(), 1e-300))
print("final SNR", snr(want, wav))
def ours(name, dtype, shape):
    return np.fromfile(os.path.join(dbg, name + ".bin"), dtype=dtype).reshape(shape).astype(np.float32)

# reference nodes are [channel][time] (time fastest) unless noted; ours are [time][channel]
def cmp(label, ref_arr, our_arr):
    print(f"{label:28s} ref{ref_arr.shape} ours{our_arr.shape} SNR {snr(ref_arr, our_arr):7.2f} dB  max|ref| {np.abs(ref_arr).max():.3g}")
cmp("vq_output", ref_node("vq_output").T, ours("lat16", np.float16, (T, 512)))
cmp("pre_conv_output", ref_node("pre_conv_output").T, ours("pc16", np.float16, (T, 1024)))
cmp("pre_tfm_input", ref_node("pre_tfm_input"), ours("tfm_in", np.float32, (T, 512)))
cmp("pre_tfm_output", ref_node("pre_tfm_output"), ours("tfm_out16", np.float16, (T, 1024)))
cmp("upsample_output", ref_node("upsample_output").T, ours("up1_16", np.float16, (4 * T, 1024)))
L = 4 * T
for i, (s, C) in enumerate(((8, 768), (5, 384), (4, 192), (3, 96)), start=1):
    L = (L - 1) * s
    # decN_output is the block output; compare our last residual-stream dump of the block instead
    try:
        cmp(f"dec{i}_output~res1", ref_node(f"dec{i}_output").T, ours(f"blk{i}_res1_32", np.float32, (L, C)))
    except Exception as e:
        print("dec", i, e)
