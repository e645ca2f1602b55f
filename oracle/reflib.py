"""TEST INFRASTRUCTURE — ctypes binding of oracle/_ref/libq3t_ref.so (the unmodified reference's
ggml-CPU path, built by oracle/Makefile.ref from /root/reference where it lies).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product (qwen3-tts.cpp_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "libq3t_ref.so")

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)


def available() -> bool:
    return os.path.exists(REF_SO)


def _fp(a):
    return None if a is None else a.ctypes.data_as(_f32p)


def _ip(a):
    return None if a is None else a.ctypes.data_as(_i32p)


class RefLib:
    _inst = None

    def __init__(self):
        L = C.CDLL(REF_SO)
        L.ref_last_error.restype = C.c_char_p
        L.ref_set_threads.argtypes = [C.c_int]
        L.ref_tts_load.restype = C.c_void_p
        L.ref_tts_load.argtypes = [C.c_char_p]
        L.ref_tts_free.argtypes = [C.c_void_p]
        L.ref_tts_config.argtypes = [C.c_void_p, _i32p, C.c_int]
        L.ref_tts_init_kv.argtypes = [C.c_void_p, C.c_int32]
        L.ref_tts_build_prompt.argtypes = [C.c_void_p, _i32p, C.c_int32, _f32p, C.c_int32,
                                           _f32p, _i32p, _f32p, _i32p, _f32p]
        L.ref_tts_prefill.argtypes = [C.c_void_p, _f32p, C.c_int32, C.c_int32, _f32p, _f32p]
        L.ref_tts_step.argtypes = [C.c_void_p, _f32p, C.c_int32, _f32p, _f32p]
        L.ref_tts_predict_codes_ar.argtypes = [C.c_void_p, _f32p, C.c_int32, _i32p, C.c_float, C.c_int32]
        L.ref_tts_predict_codes.argtypes = [C.c_void_p, _f32p, _i32p, _f32p]
        L.ref_tts_generate.argtypes = [C.c_void_p, _i32p, C.c_int32, _f32p, C.c_int32, _i32p,
                                       C.c_int32, C.c_float, C.c_float, C.c_int32]
        L.ref_tts_forward_text.argtypes = [C.c_void_p, _i32p, C.c_int32, _f32p, C.c_int32, _f32p]
        L.ref_tts_forward_codec.argtypes = [C.c_void_p, C.c_int32, C.c_int32, _f32p]
        L.ref_dec_load.restype = C.c_void_p
        L.ref_dec_load.argtypes = [C.c_char_p]
        L.ref_dec_free.argtypes = [C.c_void_p]
        L.ref_dec_decode.restype = C.c_long
        L.ref_dec_decode.argtypes = [C.c_void_p, _i32p, C.c_int32, _f32p, C.c_long]
        L.ref_enc_embed.argtypes = [C.c_char_p, _f32p, C.c_int32, _f32p, C.c_int]
        self.L = L

    @classmethod
    def get(cls) -> "RefLib":
        if cls._inst is None:
            cls._inst = RefLib()
        return cls._inst

    def err(self) -> str:
        return self.L.ref_last_error().decode()

    def set_threads(self, n: int):
        if self.L.ref_set_threads(n) != 0:
            raise RuntimeError(self.err())


class RefTTS:
    """Reference qwen3_tts::TTSTransformer (src/tts_transformer.h:190-280)."""

    H, V, CPV = 1024, 3072, 2048

    def __init__(self, path: str):
        self.lib = RefLib.get()
        self.h = self.lib.L.ref_tts_load(path.encode())
        if not self.h:
            raise RuntimeError("ref_tts_load: " + self.lib.err())
        cfg = np.zeros(15, np.int32)
        self.lib.L.ref_tts_config(self.h, _ip(cfg), 15)
        self.cfg = cfg
        self.H, self.V, self.CPV = int(cfg[2]), int(cfg[8]), int(cfg[11])

    def close(self):
        if self.h:
            self.lib.L.ref_tts_free(self.h)
            self.h = None

    def init_kv(self, n_ctx: int):
        if self.lib.L.ref_tts_init_kv(self.h, n_ctx) != 0:
            raise RuntimeError(self.lib.err())

    def build_prompt(self, tokens, spk, language_id=2050):
        tokens = np.ascontiguousarray(tokens, np.int32)
        spk = None if spk is None else np.ascontiguousarray(spk, np.float32)
        pre = np.zeros((16, self.H), np.float32)
        trail = np.zeros((len(tokens) + 2, self.H), np.float32)
        pad = np.zeros(self.H, np.float32)
        pl, tl = C.c_int32(0), C.c_int32(0)
        rc = self.lib.L.ref_tts_build_prompt(self.h, _ip(tokens), len(tokens), _fp(spk), language_id,
                                             _fp(pre), C.byref(pl), _fp(trail), C.byref(tl), _fp(pad))
        if rc != 0:
            raise RuntimeError(self.lib.err())
        return pre[:pl.value].copy(), trail[:tl.value].copy(), pad

    def prefill(self, embd, n_past=0):
        embd = np.ascontiguousarray(embd, np.float32)
        n = embd.shape[0]
        hid = np.zeros((n, self.H), np.float32)
        logits = np.zeros(self.V, np.float32)
        if self.lib.L.ref_tts_prefill(self.h, _fp(embd), n, n_past, _fp(hid), _fp(logits)) != 0:
            raise RuntimeError(self.lib.err())
        return hid, logits

    def step(self, embd, n_past):
        embd = np.ascontiguousarray(embd, np.float32)
        logits = np.zeros(self.V, np.float32)
        hid = np.zeros(self.H, np.float32)
        if self.lib.L.ref_tts_step(self.h, _fp(embd), n_past, _fp(logits), _fp(hid)) != 0:
            raise RuntimeError(self.lib.err())
        return logits, hid

    def predict_codes_ar(self, hidden, cb0, temperature=0.0, top_k=0):
        hidden = np.ascontiguousarray(hidden, np.float32)
        out = np.zeros(15, np.int32)
        if self.lib.L.ref_tts_predict_codes_ar(self.h, _fp(hidden), int(cb0), _ip(out), temperature, top_k) < 0:
            raise RuntimeError(self.lib.err())
        return out

    def predict_codes(self, hidden, prev_codes=None):
        hidden = np.ascontiguousarray(hidden, np.float32)
        prev = None if prev_codes is None else np.ascontiguousarray(prev_codes, np.int32)
        out = np.zeros((15, self.CPV), np.float32)
        if self.lib.L.ref_tts_predict_codes(self.h, _fp(hidden), _ip(prev), _fp(out)) < 0:
            raise RuntimeError(self.lib.err())
        return out

    def forward_text(self, tokens, spk, n_past=0):
        tokens = np.ascontiguousarray(tokens, np.int32)
        spk = None if spk is None else np.ascontiguousarray(spk, np.float32)
        hid = np.zeros((len(tokens), self.H), np.float32)
        if self.lib.L.ref_tts_forward_text(self.h, _ip(tokens), len(tokens), _fp(spk), n_past, _fp(hid)) < 0:
            raise RuntimeError(self.lib.err())
        return hid

    def forward_codec(self, token, n_past):
        logits = np.zeros(self.V, np.float32)
        if self.lib.L.ref_tts_forward_codec(self.h, int(token), n_past, _fp(logits)) < 0:
            raise RuntimeError(self.lib.err())
        return logits

    def generate(self, tokens, spk, max_len, language_id=2050, rep_penalty=1.05, temperature=0.0, top_k=0):
        tokens = np.ascontiguousarray(tokens, np.int32)
        spk = None if spk is None else np.ascontiguousarray(spk, np.float32)
        out = np.zeros((max(max_len, 1), 16), np.int32)
        n = self.lib.L.ref_tts_generate(self.h, _ip(tokens), len(tokens), _fp(spk), max_len, _ip(out),
                                        language_id, rep_penalty, temperature, top_k)
        if n < 0:
            raise RuntimeError(self.lib.err())
        return out[:n].copy()


class RefDecoder:
    """Reference qwen3_tts::AudioTokenizerDecoder (src/audio_tokenizer_decoder.h:159-179)."""

    def __init__(self, path: str):
        self.lib = RefLib.get()
        self.h = self.lib.L.ref_dec_load(path.encode())
        if not self.h:
            raise RuntimeError("ref_dec_load: " + self.lib.err())

    def close(self):
        if self.h:
            self.lib.L.ref_dec_free(self.h)
            self.h = None

    def decode(self, codes):
        codes = np.ascontiguousarray(codes, np.int32)
        T = codes.shape[0]
        cap = 1920 * T + 16
        out = np.zeros(cap, np.float32)
        n = self.lib.L.ref_dec_decode(self.h, _ip(codes), T, _fp(out), cap)
        if n < 0:
            raise RuntimeError(self.lib.err())
        return out[:n].copy()


def trace_decoder(dec: "RefDecoder", codes):
    """Debug aid: run one decode with ggml's eval callback and return {node name: float array (ggml order)}."""
    L = dec.lib.L
    L.ref_dec_trace.argtypes = [C.c_void_p, _i32p, C.c_int32]
    L.ref_trace_info.argtypes = [C.c_int, C.c_char_p, C.c_char_p, C.POINTER(C.c_int64), C.POINTER(C.c_int), C.POINTER(C.c_long)]
    L.ref_trace_data.argtypes = [C.c_int, _f32p]
    codes = np.ascontiguousarray(codes, np.int32)
    n = L.ref_dec_trace(dec.h, _ip(codes), codes.shape[0])
    if n < 0:
        raise RuntimeError(dec.lib.err())
    out = {}
    for i in range(n):
        name, op = C.create_string_buffer(64), C.create_string_buffer(32)
        ne = (C.c_int64 * 4)(); ty = C.c_int(0); cnt = C.c_long(0)
        L.ref_trace_info(i, name, op, ne, C.byref(ty), C.byref(cnt))
        nm = name.value.decode()
        if cnt.value == 0:
            continue
        buf = np.zeros(cnt.value, np.float32)
        L.ref_trace_data(i, _fp(buf))
        out[f"{i:04d}:{op.value.decode()}:{nm}"] = (buf, [int(x) for x in ne])
    return out


def speaker_embed(tts_model_path: str, samples_24k: np.ndarray) -> np.ndarray:
    lib = RefLib.get()
    s = np.ascontiguousarray(samples_24k, np.float32)
    out = np.zeros(1024, np.float32)
    n = lib.L.ref_enc_embed(tts_model_path.encode(), _fp(s), len(s), _fp(out), 1024)
    if n < 0:
        raise RuntimeError(lib.err())
    return out[:n]
