"""TEST INFRASTRUCTURE — Python driver of the C restatement oracle (oracle/q3t_oracle.c).

Control flow restated from the reference's TTSTransformer (src/tts_transformer.cpp):
prompt assembly :984-1122, forward_prefill/forward_step :1831-2108, predict_codes_autoregressive
:2310-2575, generate :2577-2851.  All arithmetic happens in the C library; this file only wires
tensors (memory-mapped straight from the GGUF) and the token loop.  Never imported by the product.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORC_SO = os.path.join(_HERE, "_build", "libq3t_oracle.so")

_f32p = C.POINTER(C.c_float)
_vp = C.c_void_p


def build():
    subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return ORC_SO


class OrcLayer(C.Structure):
    _fields_ = [("wtype", C.c_int)] + [(n, _vp) for n in ("wq", "wk", "wv", "wo", "wgate", "wup", "wdown")] + \
               [(n, _vp) for n in ("attn_norm", "ffn_norm", "q_norm", "k_norm")]


class OrcStack(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("n_layers", "hidden", "n_head", "n_kv_head", "head_dim", "ffn")] + \
               [("eps", C.c_float), ("rope_theta", C.c_float), ("layers", C.POINTER(OrcLayer)), ("out_norm", _vp),
                ("kcache", _vp), ("vcache", _vp), ("n_ctx", C.c_int)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORC_SO):
            build()
        L = C.CDLL(ORC_SO)
        L.orc_linear.argtypes = [_vp, C.c_int, C.c_int, C.c_int, _f32p, C.c_int, _f32p]
        L.orc_linear_wf32.argtypes = [_vp, C.c_int, C.c_int, C.c_int, _f32p, C.c_int, _f32p]
        L.orc_stack_forward.argtypes = [C.POINTER(OrcStack), _f32p, C.c_int, C.c_int, C.c_int, _f32p]
        L.orc_argmax.argtypes = [_f32p, C.c_int]
        L.orc_silu.argtypes = [_f32p, _f32p, C.c_int]
        L.orc_r16.restype = C.c_float
        L.orc_r16.argtypes = [C.c_float]
        _lib = L
    return _lib


def _fp(a):
    return a.ctypes.data_as(_f32p)


class GGUFTensors:
    """name -> (numpy view of the raw bytes, ggml type id, ggml ne)"""

    def __init__(self, path):
        import gguf
        self.reader = gguf.GGUFReader(path)
        self.t = {}
        for t in self.reader.tensors:
            self.t[t.name] = (t.data, int(t.tensor_type), [int(x) for x in t.shape])
        self.kv = {}
        for k, f in self.reader.fields.items():
            try:
                if len(f.types) == 1 and f.types[0].name in ("UINT32", "INT32", "FLOAT32", "UINT64", "BOOL", "UINT8"):
                    self.kv[k] = f.parts[f.data[0]][0].item()
            except Exception:
                pass

    def ptr(self, name):
        return self.t[name][0].ctypes.data

    def typ(self, name):
        return self.t[name][1]

    def f32(self, name):
        d, ty, ne = self.t[name]
        if ty == 0:
            return np.ascontiguousarray(d, np.float32).reshape(-1)
        if ty == 1:
            return np.ascontiguousarray(d).view(np.float16).astype(np.float32).reshape(-1)
        raise ValueError(name)


class PortTTS:
    H, NH, NKV, HD, FFN = 1024, 16, 8, 128, 3072

    def __init__(self, path):
        self.L = lib()
        self.g = GGUFTensors(path)
        kv = self.g.kv
        self.n_layers = kv.get("qwen3-tts.block_count", 28)
        self.cp_layers = kv.get("qwen3-tts.code_predictor.layer_count", 5)
        self.vocab = kv.get("qwen3-tts.vocab_size", 3072)
        self.cp_vocab = kv.get("qwen3-tts.code_predictor.vocab_size", 2048)
        self.text_dim = kv.get("qwen3-tts.text_hidden_size", 2048)
        self.eps = kv.get("qwen3-tts.attention.layer_norm_rms_epsilon", 1e-6)
        self.theta = kv.get("qwen3-tts.rope.freq_base", 1e6)
        self.ids = dict(codec_pad=kv.get("qwen3-tts.codec.pad_id", 2148), codec_bos=kv.get("qwen3-tts.codec.bos_id", 2149),
                        codec_eos=kv.get("qwen3-tts.codec.eos_id", 2150), tts_bos=kv.get("qwen3-tts.tts_bos_token_id", 151672),
                        tts_eos=kv.get("qwen3-tts.tts_eos_token_id", 151673), tts_pad=kv.get("qwen3-tts.tts_pad_token_id", 151671),
                        think=2154, nothink=2155, think_bos=2156, think_eos=2157)
        self._keep = []
        self.talker = self._stack("talker", self.n_layers, "talker.output_norm.weight", 0)
        self.cp = self._stack("code_pred", self.cp_layers, "code_pred.output_norm.weight", 16)
        self.last_hidden = None

    def _stack(self, prefix, n_layers, out_norm, n_ctx):
        g = self.g
        layers = (OrcLayer * n_layers)()
        for l in range(n_layers):
            b = f"{prefix}.blk.{l}."
            L = layers[l]
            L.wtype = g.typ(b + "attn_q.weight")
            for f, n in (("wq", "attn_q"), ("wk", "attn_k"), ("wv", "attn_v"), ("wo", "attn_output"),
                         ("wgate", "ffn_gate"), ("wup", "ffn_up"), ("wdown", "ffn_down")):
                setattr(L, f, g.ptr(b + n + ".weight"))
            for f, n in (("attn_norm", "attn_norm"), ("ffn_norm", "ffn_norm"), ("q_norm", "attn_q_norm"), ("k_norm", "attn_k_norm")):
                setattr(L, f, g.ptr(b + n + ".weight"))
        s = OrcStack()
        s.n_layers, s.hidden, s.n_head, s.n_kv_head, s.head_dim, s.ffn = n_layers, self.H, self.NH, self.NKV, self.HD, self.FFN
        s.eps, s.rope_theta = self.eps, self.theta
        s.layers = C.cast(layers, C.POINTER(OrcLayer))
        s.out_norm = g.ptr(out_norm)
        self._keep.append(layers)
        st = dict(s=s, n_layers=n_layers)
        self._alloc_kv(st, max(n_ctx, 1))
        return st

    def _alloc_kv(self, st, n_ctx):
        kc = np.zeros((st["n_layers"], n_ctx, self.NKV * self.HD), np.uint16)
        vc = np.zeros_like(kc)
        st["kc"], st["vc"] = kc, vc
        st["s"].kcache, st["s"].vcache, st["s"].n_ctx = kc.ctypes.data, vc.ctypes.data, n_ctx

    def init_kv(self, n_ctx):
        self._alloc_kv(self.talker, n_ctx)

    def _forward(self, st, x, pos0, f32_attn=0):
        x = np.ascontiguousarray(x, np.float32).reshape(-1, self.H)
        out = np.zeros_like(x)
        self.L.orc_stack_forward(C.byref(st["s"]), _fp(x), x.shape[0], pos0, f32_attn, _fp(out))
        return out

    def _linear(self, name, x, n_out, k):
        x = np.ascontiguousarray(x, np.float32).reshape(-1, k)
        y = np.zeros((x.shape[0], n_out), np.float32)
        self.L.orc_linear(self.g.ptr(name), self.g.typ(name), n_out, k, _fp(x), x.shape[0], _fp(y))
        return y

    def _rows(self, name, ids, dim):
        d = self.g.t[name][0]
        tab = np.ascontiguousarray(d).view(np.float16).reshape(-1, dim)
        return tab[np.asarray(ids, np.int64)].astype(np.float32)

    # project_text_tokens (:917-982)
    def project_text(self, tokens):
        x = self._rows("talker.text_embd.weight", tokens, self.text_dim)
        h = self._linear("talker.text_proj.fc1.weight", x, self.text_dim, self.text_dim) + self.g.f32("talker.text_proj.fc1.bias")
        h = np.ascontiguousarray(h, np.float32)
        for r in range(h.shape[0]):          # ggml_silu runs per row (vec.cpp:397-432)
            self.L.orc_silu(_fp(h[r]), _fp(h[r]), h.shape[1])
        return self._linear("talker.text_proj.fc2.weight", h, self.H, self.text_dim) + self.g.f32("talker.text_proj.fc2.bias")

    # build_prefill_graph (:984-1122)
    def build_prompt(self, tokens, spk, language_id=2050):
        i = self.ids
        sp = self.project_text([i["tts_bos"], i["tts_eos"], i["tts_pad"]])
        bos_e, eos_e, pad_e = sp[0], sp[1], sp[2]
        role = self.project_text(tokens[:3])
        cp_tok = [i["nothink"], i["think_bos"], i["think_eos"]] if language_id < 0 else [i["think"], i["think_bos"], language_id, i["think_eos"]]
        rows = [self._rows("talker.codec_embd.weight", cp_tok, self.H)]
        if spk is not None:
            rows.append(np.asarray(spk, np.float32).reshape(1, self.H))
        rows.append(self._rows("talker.codec_embd.weight", [i["codec_pad"], i["codec_bos"]], self.H))
        codec_in = np.concatenate(rows, 0)
        Lc = codec_in.shape[0]
        overlay = np.stack([(bos_e if t == Lc - 2 else pad_e) + codec_in[t] for t in range(Lc - 1)])
        first = self.project_text(tokens[3:4])[0] + codec_in[Lc - 1]
        prefill = np.concatenate([role, overlay, first[None]], 0).astype(np.float32)
        tc = max(0, len(tokens) - 9)
        trailing = np.concatenate([self.project_text(tokens[4:4 + tc]) if tc > 0 else np.zeros((0, self.H), np.float32), eos_e[None]], 0)
        return prefill, trailing.astype(np.float32), pad_e.astype(np.float32)

    def prefill(self, embd, n_past=0):
        hid = self._forward(self.talker, embd, n_past)
        self.last_hidden = hid[-1].copy()
        return hid, self._linear("talker.codec_head.weight", hid[-1], self.vocab, self.H)[0]

    def step(self, embd, n_past):
        hid = self._forward(self.talker, embd, n_past)[0]
        self.last_hidden = hid.copy()
        return self._linear("talker.codec_head.weight", hid, self.vocab, self.H)[0], hid

    # predict_codes_autoregressive, greedy (:2310-2575)
    def predict_codes_ar(self, hidden, cb0):
        cb0_e = self._rows("talker.codec_embd.weight", [cb0], self.H)[0]
        h = self._forward(self.cp, np.stack([np.asarray(hidden, np.float32), cb0_e]), 0, f32_attn=1)
        out = np.zeros(15, np.int32)
        logits = self._linear("code_pred.lm_head.0.weight", h[1], self.cp_vocab, self.H)[0]
        out[0] = int(self.L.orc_argmax(_fp(logits), self.cp_vocab))
        for s in range(1, 15):
            x = self._rows(f"code_pred.codec_embd.{s - 1}.weight", [int(out[s - 1])], self.H)[0]
            h = self._forward(self.cp, x, s + 1)[0]
            logits = self._linear(f"code_pred.lm_head.{s}.weight", h, self.cp_vocab, self.H)[0]
            out[s] = int(self.L.orc_argmax(_fp(logits), self.cp_vocab))
        return out

    # predict_codes, the legacy parallel-heads graph (:2130-2180, build_code_pred_graph :1413-1532): one token,
    # no RoPE / cache (a one-key soft-max is exactly 1, so the mixed value is the token's un-rounded V row =
    # the F32-attention numerics at position 0), no output norm, all 15 heads on the same layer output
    def predict_codes(self, hidden, prev_codes=None):
        x = np.asarray(hidden, np.float32).copy()
        if prev_codes is not None:
            for cb in range(15):
                x = x + self._rows(f"code_pred.codec_embd.{cb}.weight", [int(prev_codes[cb])], self.H)[0]
        s = self.cp["s"]
        keep = s.out_norm
        s.out_norm = None
        try:
            h = self._forward(self.cp, x, 0, f32_attn=1)[0]
        finally:
            s.out_norm = keep
        return np.stack([self._linear(f"code_pred.lm_head.{cb}.weight", h, self.cp_vocab, self.H)[0] for cb in range(15)])

    # generate, greedy (:2577-2851)
    def generate(self, tokens, spk, max_len, language_id=2050, rep_penalty=1.05):
        if max_len <= 0:
            return np.zeros((0, 16), np.int32)
        pre, trail, pad = self.build_prompt(tokens, spk, language_id)
        self.init_kv(pre.shape[0] + max_len + 8)
        _, logits = self.prefill(pre, 0)
        n_past = pre.shape[0]
        seen = set()
        frames = []
        for frame in range(max_len):
            logits = logits.copy()
            logits[self.vocab - 1024:self.ids["codec_eos"]] = -np.inf
            logits[self.ids["codec_eos"] + 1:] = -np.inf
            if rep_penalty != 1.0:
                for tok in seen:
                    logits[tok] = logits[tok] / np.float32(rep_penalty) if logits[tok] > 0 else logits[tok] * np.float32(rep_penalty)
            cb0 = int(self.L.orc_argmax(_fp(np.ascontiguousarray(logits, np.float32)), self.vocab))
            if cb0 == self.ids["codec_eos"]:
                break
            seen.add(cb0)
            rest = self.predict_codes_ar(self.last_hidden, cb0)
            frames.append([cb0] + [int(c) for c in rest])
            if frame + 1 >= max_len:
                break
            e = self._rows("talker.codec_embd.weight", [cb0], self.H)[0].copy()
            for cb in range(1, 16):
                e += self._rows(f"code_pred.codec_embd.{cb - 1}.weight", [int(rest[cb - 1])], self.H)[0]
            e += trail[frame] if frame < trail.shape[0] else pad
            logits, _ = self.step(e.astype(np.float32), n_past)
            n_past += 1
        return np.asarray(frames, np.int32).reshape(-1, 16)


# ---------------------------------------------------------------------------------------------------
# Sampling (src/tts_transformer.cpp:2349-2387 `sample_or_argmax`, and the codebook-0 variant with its
# post-processing :2664-2722).  The reference draws with std::discrete_distribution over an unseedable
# std::mt19937 (src/tts_transformer.h:336), so the oracle is the PROBABILITY VECTOR the reference hands to
# discrete_distribution, computed with the reference's own arithmetic: in-place float division by the
# temperature, k-th largest value as threshold with a `<` test (ties at the threshold all survive), float
# expf(l - max), double-precision sum, (float)(e / sum).
def sample_probs(logits, temperature, top_k, process=False, seen=(), rep_penalty=1.0, suppress_start=2048, eos_id=2150):
    l = np.array(logits, np.float32, copy=True)
    if process:                                       # :2666-2683 (codebook 0 only)
        for i in range(suppress_start, len(l)):
            if i != eos_id:
                l[i] = -np.inf
        if rep_penalty != 1.0:
            rp = np.float32(rep_penalty)
            for tok in sorted(set(int(t) for t in seen)):
                if 0 <= tok < len(l):
                    l[tok] = l[tok] / rp if l[tok] > 0 else l[tok] * rp
    l = (l / np.float32(temperature)).astype(np.float32)
    if 0 < top_k < len(l):
        thr = np.sort(l)[::-1][top_k - 1]             # partial_sort ... scored[top_k - 1].first
        l[l < thr] = -np.inf
    mx = l.max()
    with np.errstate(over="ignore", under="ignore"):
        e = np.exp((l - mx).astype(np.float32)).astype(np.float32)      # expf per element (float result)
    total = float(np.sum(e.astype(np.float64)))
    return (e.astype(np.float64) / total).astype(np.float32), l
