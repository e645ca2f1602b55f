"""ctypes binding of include/q3t.h (host-class C view + a few engine entry points)."""
from __future__ import annotations

import ctypes as C
import os
import re
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libq3t_b200.so")
HEADER = os.path.join(os.path.dirname(_HERE), "include", "q3t.h")

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)
_i64p = C.POINTER(C.c_int64)


class Q3TError(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile libq3t_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", _HERE, "-j8"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:], r.stderr[-4000:])
    if r.returncode != 0:
        raise Q3TError("building libq3t_b200.so failed")
    return LIB_PATH


def declared_symbols() -> list[str]:
    """Every function include/q3t.h declares with Q3T_API."""
    src = open(HEADER).read()
    return sorted(set(re.findall(r"Q3T_API[^;(]*?\b(q3t_[a-z0-9_]+)\s*\(", src)))


class TTSConfig(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("text_vocab_size", "text_embd_dim", "hidden_size", "n_layers",
                                          "n_attention_heads", "n_key_value_heads", "intermediate_size", "head_dim")] + \
               [("rms_norm_eps", C.c_float), ("rope_theta", C.c_float)] + \
               [(n, C.c_int32) for n in ("codec_vocab_size", "n_codebooks", "code_pred_layers", "code_pred_vocab_size",
                                          "codec_pad_id", "codec_bos_id", "codec_eos_id", "tts_bos_token_id",
                                          "tts_eos_token_id", "tts_pad_token_id", "codec_think_id", "codec_nothink_id",
                                          "codec_think_bos_id", "codec_think_eos_id", "english_language_id")]


class StreamStats(C.Structure):
    _fields_ = [("t_first_audio_ms", C.c_double), ("t_total_ms", C.c_double), ("n_frames", C.c_int32), ("n_chunks", C.c_int32),
                ("n_samples", C.c_int64)]


class BatchStats(C.Structure):
    _fields_ = [("t_generate_ms", C.c_double), ("t_decode_ms", C.c_double), ("t_total_ms", C.c_double), ("n_frames", C.c_int64),
                ("n_samples", C.c_int64)]


class Timing(C.Structure):
    _fields_ = [("prefill_ms", C.c_double), ("frames_ms", C.c_double), ("vocoder_ms", C.c_double),
                ("n_frames", C.c_int32), ("kernel_launches", C.c_int64)]


_lib = None


def lib() -> C.CDLL:
    """Load the product library; fail loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Q3TError(f"{LIB_PATH} is missing: run `make -C qwen3-tts.cpp_b200` (or __graft_entry__.build()); "
                       "there is no fallback implementation")
    L = C.CDLL(LIB_PATH)
    vp, cp = C.c_void_p, C.c_char_p
    sig = {
        "q3t_last_error": (cp, [vp]),
        "q3t_gguf_open": (vp, [cp]), "q3t_gguf_close": (None, [vp]), "q3t_gguf_get_error": (cp, [vp]),
        "q3t_gguf_get_n_tensors": (C.c_int64, [vp]), "q3t_gguf_get_tensor_name": (cp, [vp, C.c_int64]),
        "q3t_gguf_get_tensor_type": (C.c_int32, [vp, C.c_int64]), "q3t_gguf_get_tensor_offset": (C.c_size_t, [vp, C.c_int64]),
        "q3t_gguf_get_tensor_size": (C.c_size_t, [vp, C.c_int64]), "q3t_gguf_get_tensor_ne": (C.c_int32, [vp, C.c_int64, _i64p]),
        "q3t_gguf_get_u32": (C.c_int32, [vp, cp, C.c_int32]), "q3t_gguf_get_f32": (C.c_float, [vp, cp, C.c_float]),
        "q3t_gguf_get_data_offset": (C.c_size_t, [vp]),
        "q3t_tts_new": (vp, []), "q3t_tts_delete": (None, [vp]), "q3t_tts_load_model": (C.c_int, [vp, cp]),
        "q3t_tts_unload_model": (None, [vp]), "q3t_tts_init_kv_cache": (C.c_int, [vp, C.c_int32]),
        "q3t_tts_clear_kv_cache": (None, [vp]),
        "q3t_tts_forward_prefill": (C.c_int, [vp, _f32p, C.c_int32, C.c_int32, _f32p, _f32p]),
        "q3t_tts_forward_step": (C.c_int, [vp, _f32p, C.c_int32, _f32p, _f32p]),
        "q3t_tts_forward_text": (C.c_int, [vp, _i32p, C.c_int32, _f32p, C.c_int32, _f32p]),
        "q3t_tts_forward_codec": (C.c_int, [vp, C.c_int32, C.c_int32, _f32p]),
        "q3t_tts_get_hidden_states": (C.c_int, [vp, _f32p]),
        "q3t_tts_predict_codes": (C.c_int, [vp, _f32p, _i32p, _f32p]),
        "q3t_tts_predict_codes_autoregressive": (C.c_int, [vp, _f32p, C.c_int32, _i32p, C.c_float, C.c_int32]),
        "q3t_tts_generate": (C.c_int, [vp, _i32p, C.c_int32, _f32p, C.c_int32, _i32p, _i32p, C.c_int32, C.c_float, C.c_float, C.c_int32]),
        "q3t_tts_generate_batch": (C.c_int, [vp, C.c_int32, C.POINTER(_i32p), _i32p, C.POINTER(_f32p), C.c_int32, C.POINTER(_i32p), _i32p, C.c_int32, C.c_float]),
        "q3t_tts_generate_batch_sampled": (C.c_int, [vp, C.c_int32, C.POINTER(_i32p), _i32p, C.POINTER(_f32p), C.c_int32, C.POINTER(_i32p), _i32p, C.c_int32, C.c_float, C.c_float, C.c_int32]),
        "q3t_tts_set_seed": (None, [vp, C.c_uint64]), "q3t_tts_get_config": (C.c_int, [vp, C.POINTER(TTSConfig)]),
        "q3t_tts_get_error": (cp, [vp]), "q3t_tts_engine": (vp, [vp]),
        "q3t_dec_new": (vp, []), "q3t_dec_delete": (None, [vp]), "q3t_dec_load_model": (C.c_int, [vp, cp]),
        "q3t_dec_unload_model": (None, [vp]), "q3t_dec_decode": (C.c_int64, [vp, _i32p, C.c_int32, _f32p, C.c_int64]),
        "q3t_dec_get_error": (cp, [vp]), "q3t_dec_engine": (vp, [vp]),
        "q3t_dec_stream_reset": (None, [vp]), "q3t_dec_decode_append": (C.c_int64, [vp, _i32p, C.c_int32, _f32p, C.c_int64]),
        "q3t_build_prompt": (C.c_int, [vp, _i32p, C.c_int32, _f32p, C.c_int32, _f32p, _i32p, _f32p, _i32p, _f32p]),
        "q3t_project_text": (C.c_int, [vp, _i32p, C.c_int32, _f32p]),
        "q3t_enc_new": (vp, []), "q3t_enc_delete": (None, [vp]), "q3t_enc_load_model": (C.c_int, [vp, cp]),
        "q3t_enc_encode": (C.c_int, [vp, _f32p, C.c_int32, _f32p]), "q3t_enc_get_error": (cp, [vp]),
        "q3t_speaker_encode": (C.c_int, [vp, _f32p, C.c_int32, _f32p, _i32p]),
        "q3t_stream_begin": (C.c_int, [vp, _i32p, C.c_int32, _f32p, C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_int32, C.c_uint64]),
        "q3t_stream_next": (C.c_int, [vp, C.c_int32, _i32p, _i32p, _i32p]),
        "q3t_tts_generate_begin": (C.c_int, [vp, _i32p, C.c_int32, _f32p, C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_int32]),
        "q3t_tts_generate_next": (C.c_int, [vp, C.c_int32, _i32p, _i32p, _i32p]),
        "q3t_pipe_synthesize_stream": (C.c_int64, [vp, vp, _i32p, C.c_int32, _f32p, C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_int32,
                                                   C.c_int32, C.c_int32, _f32p, C.c_int64, C.POINTER(StreamStats)]),
        "q3t_pipe_synthesize_batch": (C.c_int, [vp, vp, C.c_int32, C.POINTER(_i32p), _i32p, C.POINTER(_f32p), C.c_int32, C.c_int32, C.c_float,
                                                C.POINTER(_f32p), _i64p, _i64p, C.POINTER(BatchStats)]),
        "q3t_pipe_synthesize_batch_multi": (C.c_int, [vp, C.POINTER(vp), C.c_int32, C.c_int32, C.POINTER(_i32p), _i32p, C.POINTER(_f32p), C.c_int32, C.c_int32,
                                                      C.c_float, C.POINTER(_f32p), _i64p, _i64p, C.POINTER(BatchStats)]),
        "q3t_pipe_decode_batch_multi": (C.c_int, [C.POINTER(vp), C.c_int32, C.c_int32, C.POINTER(_i32p), _i32p, C.POINTER(_f32p), _i64p, _i64p]),
        "q3t_pipe_get_error": (cp, []),
        "q3t_sample_logits": (C.c_int, [vp, _f32p, C.c_int32, C.c_int32, _i32p, C.c_int32, C.c_float, C.c_float, C.c_int32, C.c_uint64, C.c_int32, _i32p]),
        "q3t_project_text_static_dev": (C.c_int, [vp, vp, C.c_int32, vp]),
        "q3t_project_text_static_max_tokens": (C.c_int, []),
        "q3t_stage_prompt": (C.c_int, [vp, _i32p, C.c_int32, _f32p, C.c_int32]),
        "q3t_run_staged": (C.c_int, [vp, C.c_int32, C.c_float, C.c_float, C.c_int32, C.c_uint64, C.c_int32, _i32p, _i64p]),
        "q3t_fetch_codes": (C.c_int, [vp, _i32p, C.c_int32]), "q3t_fetch_samples": (C.c_int64, [vp, _f32p, C.c_int64]),
        "q3t_bench_talker_step": (C.c_int, [vp, C.c_int32, C.c_int32, _f32p]),
        "q3t_get_timing": (C.c_int, [vp, C.POINTER(Timing)]), "q3t_weight_bytes": (C.c_int64, [vp, C.c_int32]),
        "q3t_vocoder_n_samples": (C.c_int64, [C.c_int32]), "q3t_kv_n_ctx": (C.c_int, [vp]),
        "q3t_device_info": (C.c_int, [vp, _i32p, _i32p, _i32p]),
        "q3t_profile_arm": (C.c_int, [vp, C.c_int32, C.c_int32]),
        "q3t_profile_read": (C.c_int, [vp, C.POINTER(C.c_uint64), C.c_int32]),
        "q3t_profile_select": (C.c_int, [vp, C.c_int32]),
    }
    for name, (res, args) in sig.items():
        if os.environ.get("Q3T_AB_LENIENT") and not hasattr(L, name):       # developer A/B runs against an older library build
            continue
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _fp(a):
    return None if a is None else a.ctypes.data_as(_f32p)


def _ip(a):
    return None if a is None else a.ctypes.data_as(_i32p)


class GGUFLoader:
    """qwen3_tts::GGUFLoader (src/gguf_loader.h:16-62)."""

    def __init__(self):
        self.h = None
        self._err = ""

    def open(self, path: str) -> bool:
        self.close()
        self.h = lib().q3t_gguf_open(path.encode())
        self._err = lib().q3t_gguf_get_error(self.h).decode()
        return self._err == ""

    def close(self):
        if self.h:
            lib().q3t_gguf_close(self.h)
            self.h = None

    def get_error(self) -> str:
        return self._err

    def get_n_tensors(self) -> int:
        return lib().q3t_gguf_get_n_tensors(self.h)

    def get_tensor_name(self, i: int) -> str:
        return lib().q3t_gguf_get_tensor_name(self.h, i).decode()

    def get_tensor_type(self, i: int) -> int:
        return lib().q3t_gguf_get_tensor_type(self.h, i)

    def get_tensor_offset(self, i: int) -> int:
        return lib().q3t_gguf_get_tensor_offset(self.h, i)

    def get_tensor_size(self, i: int) -> int:
        return lib().q3t_gguf_get_tensor_size(self.h, i)

    def get_tensor_ne(self, i: int):
        ne = (C.c_int64 * 4)()
        lib().q3t_gguf_get_tensor_ne(self.h, i, ne)
        return list(ne)

    def get_u32(self, key: str, default: int = 0) -> int:
        return lib().q3t_gguf_get_u32(self.h, key.encode(), default)

    def get_f32(self, key: str, default: float = 0.0) -> float:
        return lib().q3t_gguf_get_f32(self.h, key.encode(), default)

    def get_data_offset(self) -> int:
        return lib().q3t_gguf_get_data_offset(self.h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class TTSTransformer:
    """qwen3_tts::TTSTransformer (src/tts_transformer.h:190-280); methods return False and set
    get_error() on failure exactly like the reference — no exceptions on the compute path."""

    def __init__(self):
        self.L = lib()
        self.h = self.L.q3t_tts_new()
        self.cfg = None

    def close(self):
        if self.h:
            self.L.q3t_tts_delete(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def get_error(self) -> str:
        return self.L.q3t_tts_get_error(self.h).decode()

    def load_model(self, path: str) -> bool:
        ok = bool(self.L.q3t_tts_load_model(self.h, path.encode()))
        if ok:
            c = TTSConfig()
            self.L.q3t_tts_get_config(self.h, C.byref(c))
            self.cfg = c
        return ok

    def unload_model(self):
        self.L.q3t_tts_unload_model(self.h)

    def get_config(self):
        return self.cfg

    def engine(self):
        return self.L.q3t_tts_engine(self.h)

    def set_seed(self, seed: int):
        self.L.q3t_tts_set_seed(self.h, seed)

    def init_kv_cache(self, n_ctx: int) -> bool:
        return bool(self.L.q3t_tts_init_kv_cache(self.h, n_ctx))

    def clear_kv_cache(self):
        self.L.q3t_tts_clear_kv_cache(self.h)

    def forward_prefill(self, embd, n_past=0):
        embd = np.ascontiguousarray(embd, np.float32)
        n = embd.shape[0]
        hid = np.zeros((n, self.cfg.hidden_size), np.float32)
        logits = np.zeros(self.cfg.codec_vocab_size, np.float32)
        if not self.L.q3t_tts_forward_prefill(self.h, _fp(embd), n, n_past, _fp(hid), _fp(logits)):
            return None
        return hid, logits

    def forward_step(self, embd, n_past):
        embd = np.ascontiguousarray(embd, np.float32)
        logits = np.zeros(self.cfg.codec_vocab_size, np.float32)
        hid = np.zeros(self.cfg.hidden_size, np.float32)
        if not self.L.q3t_tts_forward_step(self.h, _fp(embd), n_past, _fp(logits), _fp(hid)):
            return None
        return logits, hid

    def forward_text(self, tokens, spk, n_past=0):
        tokens = np.ascontiguousarray(tokens, np.int32)
        spk = None if spk is None else np.ascontiguousarray(spk, np.float32)
        hid = np.zeros((len(tokens), self.cfg.hidden_size), np.float32)
        if not self.L.q3t_tts_forward_text(self.h, _ip(tokens), len(tokens), _fp(spk), n_past, _fp(hid)):
            return None
        return hid

    def forward_codec(self, token, n_past):
        logits = np.zeros(self.cfg.codec_vocab_size, np.float32)
        if not self.L.q3t_tts_forward_codec(self.h, int(token), n_past, _fp(logits)):
            return None
        return logits

    def get_hidden_states(self):
        hid = np.zeros(self.cfg.hidden_size, np.float32)
        if not self.L.q3t_tts_get_hidden_states(self.h, _fp(hid)):
            return None
        return hid

    def predict_codes(self, hidden, prev_codes=None):
        hidden = np.ascontiguousarray(hidden, np.float32)
        prev = None if prev_codes is None else np.ascontiguousarray(prev_codes, np.int32)
        out = np.zeros((15, self.cfg.code_pred_vocab_size if self.cfg is not None else 2048), np.float32)
        if not self.L.q3t_tts_predict_codes(self.h, _fp(hidden), _ip(prev), _fp(out)):
            return None
        return out

    def predict_codes_autoregressive(self, hidden, cb0, temperature=0.9, top_k=50):
        hidden = np.ascontiguousarray(hidden, np.float32)
        out = np.zeros(15, np.int32)
        if not self.L.q3t_tts_predict_codes_autoregressive(self.h, _fp(hidden), int(cb0), _ip(out), temperature, top_k):
            return None
        return out

    def generate(self, tokens, spk, max_len, language_id=2050, repetition_penalty=1.05, temperature=0.9, top_k=50):
        tokens = None if tokens is None else np.ascontiguousarray(tokens, np.int32)
        spk = None if spk is None else np.ascontiguousarray(spk, np.float32)
        out = np.zeros((max(max_len, 1), 16), np.int32)
        nf = C.c_int32(0)
        n_tok = 0 if tokens is None else len(tokens)
        if not self.L.q3t_tts_generate(self.h, _ip(tokens), n_tok, _fp(spk), max_len, _ip(out), C.byref(nf),
                                       language_id, repetition_penalty, temperature, top_k):
            return None
        return out[:nf.value].copy()

    def generate_stream(self, tokens, spk, max_len, chunk, language_id=2050, repetition_penalty=1.05, temperature=0.9, top_k=50):
        """generate() in chunks of `chunk` frames (generate_begin / generate_next): yields int32[n, 16] arrays."""
        tokens = np.ascontiguousarray(tokens, np.int32)
        spk = None if spk is None else np.ascontiguousarray(spk, np.float32)
        if not self.L.q3t_tts_generate_begin(self.h, _ip(tokens), len(tokens), _fp(spk), max_len, language_id,
                                             repetition_penalty, temperature, top_k):
            raise Q3TError(self.get_error())
        done = max_len <= 0
        while not done:
            buf = np.zeros((chunk, 16), np.int32)
            n, d = C.c_int32(0), C.c_int32(1)
            if not self.L.q3t_tts_generate_next(self.h, chunk, _ip(buf), C.byref(n), C.byref(d)):
                raise Q3TError(self.get_error())
            done = bool(d.value)
            if n.value:
                yield buf[:n.value].copy()

    def generate_batch(self, utterances, max_len, language_id=2050, repetition_penalty=1.05, temperature=0.0, top_k=0):
        """utterances: list of (tokens int32[], speaker f32[1024] or None); greedy unless temperature > 0 (utterance i then
        draws from the stream seeded set_seed()'s seed + i); returns a list of code arrays."""
        n = len(utterances)
        toks = [np.ascontiguousarray(u[0], np.int32) for u in utterances]
        spks = [None if u[1] is None else np.ascontiguousarray(u[1], np.float32) for u in utterances]
        outs = [np.zeros((max(max_len, 1), 16), np.int32) for _ in range(n)]
        tp = (_i32p * n)(*[_ip(t) for t in toks])
        sp = (_f32p * n)(*[_fp(x) if x is not None else C.cast(None, _f32p) for x in spks])
        op = (_i32p * n)(*[_ip(o) for o in outs])
        nt = np.array([len(t) for t in toks], np.int32)
        nf = np.zeros(n, np.int32)
        if not self.L.q3t_tts_generate_batch_sampled(self.h, n, tp, _ip(nt), sp, max_len, op, _ip(nf), language_id, repetition_penalty,
                                                     temperature, top_k):
            return None
        return [outs[i][:nf[i]].copy() for i in range(n)]

    # ---- engine-level helpers (device-resident path used by bench.py `value`)
    def build_prompt(self, tokens, spk, language_id=2050):
        tokens = np.ascontiguousarray(tokens, np.int32)
        spk = None if spk is None else np.ascontiguousarray(spk, np.float32)
        pre = np.zeros((16, self.cfg.hidden_size), np.float32)
        trail = np.zeros((len(tokens) + 2, self.cfg.hidden_size), np.float32)
        pad = np.zeros(self.cfg.hidden_size, np.float32)
        pl, tl = C.c_int32(0), C.c_int32(0)
        rc = self.L.q3t_build_prompt(self.engine(), _ip(tokens), len(tokens), _fp(spk), language_id,
                                     _fp(pre), C.byref(pl), _fp(trail), C.byref(tl), _fp(pad))
        if rc != 0:
            raise Q3TError(self.L.q3t_last_error(self.engine()).decode())
        return pre[:pl.value].copy(), trail[:tl.value].copy(), pad

    def stage_prompt(self, tokens, spk, language_id=2050):
        tokens = np.ascontiguousarray(tokens, np.int32)
        spk = None if spk is None else np.ascontiguousarray(spk, np.float32)
        rc = self.L.q3t_stage_prompt(self.engine(), _ip(tokens), len(tokens), _fp(spk), language_id)
        if rc != 0:
            raise Q3TError(self.L.q3t_last_error(self.engine()).decode())

    def run_staged(self, max_len, repetition_penalty=1.05, temperature=0.0, top_k=0, seed=0, with_vocoder=True):
        nf, ns = C.c_int32(0), C.c_int64(0)
        rc = self.L.q3t_run_staged(self.engine(), max_len, repetition_penalty, temperature, top_k, seed,
                                   1 if with_vocoder else 0, C.byref(nf), C.byref(ns))
        if rc != 0:
            raise Q3TError(self.L.q3t_last_error(self.engine()).decode())
        return nf.value, ns.value

    def fetch_codes(self, n_frames):
        out = np.zeros((n_frames, 16), np.int32)
        if self.L.q3t_fetch_codes(self.engine(), _ip(out), n_frames) != 0:
            raise Q3TError(self.L.q3t_last_error(self.engine()).decode())
        return out

    def fetch_samples(self, n):
        out = np.zeros(n, np.float32)
        got = self.L.q3t_fetch_samples(self.engine(), _fp(out), n)
        if got < 0:
            raise Q3TError(self.L.q3t_last_error(self.engine()).decode())
        return out[:got]

    def bench_talker_step(self, n_past, iters=64):
        us = C.c_float(0)
        rc = self.L.q3t_bench_talker_step(self.engine(), n_past, iters, C.byref(us))
        if rc != 0:
            raise Q3TError(self.L.q3t_last_error(self.engine()).decode())
        return us.value

    def profile_arm(self, cta=0, cap=1 << 16):
        if self.L.q3t_profile_arm(self.engine(), cta, cap) != 0:
            raise Q3TError("profile_arm failed")

    def profile_read(self, cap=1 << 16, cta=None):
        if cta is not None:
            self.L.q3t_profile_select(self.engine(), cta)
        buf = np.zeros(cap, np.uint64)
        n = self.L.q3t_profile_read(self.engine(), buf.ctypes.data_as(C.POINTER(C.c_uint64)), cap)
        if n < 0:
            raise Q3TError("profile_read failed")
        return (buf[:n] >> np.uint64(8)).astype(np.int64), (buf[:n] & np.uint64(0xff)).astype(np.int64)

    def sample_logits(self, logits, n_draws, temperature, top_k, seed=0, process=False, seen=None, repetition_penalty=1.0):
        """n_draws tokens from one logit vector through the device sampler (q3t_sample_logits)."""
        logits = np.ascontiguousarray(logits, np.float32)
        seen = np.zeros(0, np.int32) if seen is None else np.ascontiguousarray(seen, np.int32)
        out = np.zeros(n_draws, np.int32)
        rc = self.L.q3t_sample_logits(self.engine(), _fp(logits), len(logits), 1 if process else 0, _ip(seen), len(seen),
                                      repetition_penalty, temperature, top_k, seed, n_draws, _ip(out))
        if rc != 0:
            raise Q3TError(self.L.q3t_last_error(self.engine()).decode())
        return out

    def timing(self):
        t = Timing()
        self.L.q3t_get_timing(self.engine(), C.byref(t))
        return t

    def weight_bytes(self, which):
        return self.L.q3t_weight_bytes(self.engine(), which)


class AudioTokenizerDecoder:
    """qwen3_tts::AudioTokenizerDecoder (src/audio_tokenizer_decoder.h:159-179)."""

    def __init__(self):
        self.L = lib()
        self.h = self.L.q3t_dec_new()

    def close(self):
        if self.h:
            self.L.q3t_dec_delete(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def get_error(self) -> str:
        return self.L.q3t_dec_get_error(self.h).decode()

    def load_model(self, path: str) -> bool:
        return bool(self.L.q3t_dec_load_model(self.h, path.encode()))

    def unload_model(self):
        self.L.q3t_dec_unload_model(self.h)

    def decode(self, codes):
        codes = np.ascontiguousarray(codes, np.int32)
        T = codes.shape[0]
        cap = max(int(self.L.q3t_vocoder_n_samples(T)), 0) + 16
        out = np.empty(cap, np.float32)                       # written by the D2H copy: no zero fill, no second pass
        n = self.L.q3t_dec_decode(self.h, _ip(codes), T, _fp(out), cap)
        if n < 0:
            return None
        return out[:n]

    # streaming extension: the concatenation of decode_append() results equals decode() of all the frames
    def stream_reset(self):
        self.L.q3t_dec_stream_reset(self.h)

    def decode_append(self, codes):
        codes = np.ascontiguousarray(codes, np.int32)
        T = codes.shape[0]
        cap = 1920 * T + 16
        out = np.empty(cap, np.float32)
        n = self.L.q3t_dec_decode_append(self.h, _ip(codes), T, _fp(out), cap)
        if n < 0:
            return None
        return out[:n]


class AudioTokenizerEncoder:
    """qwen3_tts::AudioTokenizerEncoder (speaker x-vector; src/audio_tokenizer_encoder.h:104-150)."""

    def __init__(self):
        self.L = lib()
        self.h = self.L.q3t_enc_new()

    def close(self):
        if self.h:
            self.L.q3t_enc_delete(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def get_error(self) -> str:
        return self.L.q3t_enc_get_error(self.h).decode()

    def load_model(self, path: str) -> bool:
        return bool(self.L.q3t_enc_load_model(self.h, path.encode()))

    def encode(self, samples):
        samples = np.ascontiguousarray(samples, np.float32)
        out = np.zeros(1024, np.float32)
        if not self.L.q3t_enc_encode(self.h, _fp(samples), len(samples), _fp(out)):
            return None
        return out


def synthesize_stream(tts: TTSTransformer, dec: AudioTokenizerDecoder, tokens, spk, max_len, first_chunk=8, chunk=32,
                      language_id=2050, repetition_penalty=1.05, temperature=0.0, top_k=0):
    """Streaming front door (host/b200_pipeline.h): returns (waveform, StreamStats)."""
    tokens = np.ascontiguousarray(tokens, np.int32)
    spk = None if spk is None else np.ascontiguousarray(spk, np.float32)
    cap = 1920 * max(max_len, 1) + 16
    out = np.zeros(cap, np.float32)
    st = StreamStats()
    n = lib().q3t_pipe_synthesize_stream(tts.h, dec.h, _ip(tokens), len(tokens), _fp(spk), max_len, language_id, repetition_penalty,
                                         temperature, top_k, first_chunk, chunk, _fp(out), cap, C.byref(st))
    if n < 0:
        raise Q3TError(lib().q3t_pipe_get_error().decode())
    return out[:n].copy(), st


def decode_batch(decs, codes_list):
    """codes_list: list of int32 [T_i, 16] arrays; utterance i is decoded by decs[i mod N] (one host thread per decoder, in C++)."""
    decs = list(decs) if isinstance(decs, (list, tuple)) else [decs]
    n = len(codes_list)
    cs = [np.ascontiguousarray(c, np.int32) for c in codes_list]
    nf = np.array([len(c) for c in cs], np.int32)
    outs = [np.empty(max(1920 * int(k) - 555, 1), np.float32) for k in nf]
    caps = np.array([len(o) for o in outs], np.int64)
    ns = np.zeros(n, np.int64)
    cp_ = (_i32p * n)(*[_ip(c) for c in cs])
    op = (_f32p * n)(*[_fp(o) for o in outs])
    dp = (C.c_void_p * len(decs))(*[x.h for x in decs])
    if not lib().q3t_pipe_decode_batch_multi(dp, len(decs), n, cp_, _ip(nf), op, caps.ctypes.data_as(_i64p), ns.ctypes.data_as(_i64p)):
        raise Q3TError(lib().q3t_pipe_get_error().decode())
    return [outs[i][:ns[i]] for i in range(n)]


def synthesize_batch(tts: TTSTransformer, dec, utterances, max_len, language_id=2050, repetition_penalty=1.05):
    """Batched front door: utterances = [(tokens, spk or None)], greedy; returns ([waveform], BatchStats).
    dec: one AudioTokenizerDecoder or a list of them (utterance i is decoded by decoder i mod N, one host thread each)."""
    decs = list(dec) if isinstance(dec, (list, tuple)) else [dec]
    n = len(utterances)
    toks = [np.ascontiguousarray(u[0], np.int32) for u in utterances]
    spks = [None if u[1] is None else np.ascontiguousarray(u[1], np.float32) for u in utterances]
    cap = 1920 * max(max_len, 1) + 16
    outs = [np.zeros(cap, np.float32) for _ in range(n)]
    tp = (_i32p * n)(*[_ip(t) for t in toks])
    sp = (_f32p * n)(*[_fp(x) if x is not None else C.cast(None, _f32p) for x in spks])
    op = (_f32p * n)(*[_fp(o) for o in outs])
    nt = np.array([len(t) for t in toks], np.int32)
    caps = np.full(n, cap, np.int64)
    ns = np.zeros(n, np.int64)
    st = BatchStats()
    dp = (C.c_void_p * len(decs))(*[x.h for x in decs])
    ok = lib().q3t_pipe_synthesize_batch_multi(tts.h, dp, len(decs), n, tp, _ip(nt), sp, max_len, language_id, repetition_penalty, op,
                                               caps.ctypes.data_as(_i64p), ns.ctypes.data_as(_i64p), C.byref(st))
    if not ok:
        raise Q3TError(lib().q3t_pipe_get_error().decode())
    return [outs[i][:ns[i]].copy() for i in range(n)], st
