#!/usr/bin/env python3
"""Random-init GGUF writer for the Qwen3-TTS-12Hz-0.6B architecture in the reference's
on-disk layout (reference converters
scripts/convert_tts_to_gguf.py:244-319,440-471 and scripts/convert_tokenizer_to_gguf.py:45-166,270-300).

The real checkpoint is not available offline, so parity and benchmarks run on seeded random
weights of the real shapes.  numpy's PCG64 stream is platform independent, so the same seed
gives byte-identical files in the build container and on the GPU box.

Presets
  full : the real 0.6B shapes (28 talker layers, 5 code-predictor layers, text vocab 151936)
  tiny : same per-layer shapes (the kernels are specialised on them) but 2 talker layers,
         2 code-predictor layers and a 512-entry text vocabulary -> ~0.2 GB, fast CPU tests.
"""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np
import gguf

PRESETS = {
    "full": dict(n_layers=28, cp_layers=5, text_vocab=151936,
                 tts_bos=151672, tts_eos=151673, tts_pad=151671),
    "tiny": dict(n_layers=2, cp_layers=2, text_vocab=512,
                 tts_bos=510, tts_eos=511, tts_pad=509),
}

H, NH, NKV, HD, FFN = 1024, 16, 8, 128, 3072
TEXT_DIM = 2048
CODEC_VOCAB, CP_VOCAB, N_CB = 3072, 2048, 16


class _Rng:
    """float32 normal stream; one Generator so tensor order defines the values."""

    def __init__(self, seed: int):
        self.g = np.random.Generator(np.random.PCG64(seed))

    def normal(self, shape, std):
        a = self.g.standard_normal(size=shape, dtype=np.float32)
        a *= np.float32(std)
        return a


def _q4_k(data: np.ndarray) -> np.ndarray:
    """Q4_K blocks from ggml's own quantiser (gguf-py has none: the reference's converter falls back to F16 with this
    gguf version).  Uses ggml_quantize_chunk of the reference build under oracle/_ref -- test infrastructure only."""
    import ctypes as C
    so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libq3t_ref.so")
    L = C.CDLL(so)
    L.ggml_quantize_chunk.restype = C.c_size_t
    L.ggml_quantize_chunk.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
    rows, k = data.shape
    assert k % 256 == 0
    src = np.ascontiguousarray(data, np.float32)
    dst = np.zeros((rows, k // 256 * 144), np.uint8)
    n = L.ggml_quantize_chunk(12, src.ctypes.data, dst.ctypes.data, 0, rows, k, None)
    assert n == dst.nbytes, (n, dst.nbytes)
    return dst


def _add(writer, name, data, qtype):
    """1-D -> F32; >=2-D -> F16, F32, Q8_0 or Q4_K (quantised along the last numpy axis = ggml ne[0])."""
    if data.ndim <= 1:
        writer.add_tensor(name, data.astype(np.float32), raw_dtype=gguf.GGMLQuantizationType.F32)
    elif qtype == "q8_0":
        q = gguf.quants.quantize(data.astype(np.float32), gguf.GGMLQuantizationType.Q8_0)
        writer.add_tensor(name, q, raw_dtype=gguf.GGMLQuantizationType.Q8_0)
    elif qtype == "q4_k":
        writer.add_tensor(name, _q4_k(data.astype(np.float32)), raw_dtype=gguf.GGMLQuantizationType.Q4_K)
    elif qtype == "f32":
        writer.add_tensor(name, data.astype(np.float32), raw_dtype=gguf.GGMLQuantizationType.F32)
    else:
        writer.add_tensor(name, data.astype(np.float16), raw_dtype=gguf.GGMLQuantizationType.F16)


def gpt2_byte_chars() -> list[str]:
    """bytes_to_unicode of GPT-2 in vocabulary order: printable bytes first, the other 68 as chr(256 + n)."""
    bs = list(range(ord("!"), ord("~") + 1)) + list(range(0xA1, 0xAC + 1)) + list(range(0xAE, 0xFF + 1))
    out = [chr(b) for b in bs]
    n = 0
    for b in range(256):
        if b not in bs:
            out.append(chr(256 + n))
            n += 1
    return out


def synthetic_vocab(n: int) -> list[str]:
    toks = gpt2_byte_chars() + [f"[unused{i}]" for i in range(256, n)]
    toks[77091] = "assistant"
    toks[151643] = "<|endoftext|>"
    toks[151644] = "<|im_start|>"
    toks[151645] = "<|im_end|>"
    return toks


def write_tts_gguf(path: str, preset: str = "full", seed: int = 1234, wtype: str = "f16",
                   w_std: float = 0.02, eos_like: int = -1) -> dict:
    """Talker + code predictor file (reference name: qwen3-tts-0.6b-f16.gguf).

    wtype "q8_0" quantises exactly the tensors the reference converter quantises
    (attn q/k/v/o, ffn gate/up/down, text_proj fc1/fc2); embeddings and heads stay F16
    (scripts/convert_tts_to_gguf.py:244-270).

    eos_like >= 0 makes the EOS row (id 2150) of talker.codec_head 1.02 x the row of token `eos_like`: with plain
    random weights the EOS logit is persistently negative (the hidden state of a random model has a dominant
    direction) and never wins; as a slightly louder copy of a token that greedy decoding does emit, EOS wins at
    the first frame where that token would have been chosen, a few to a few dozen frames in and at a different
    frame per utterance -- what the early-stop tests need (every other tensor is byte-identical to the plain file).
    """
    p = PRESETS[preset]
    rng = _Rng(seed)
    arch = "qwen3-tts"
    w = gguf.GGUFWriter(path=None, arch=arch)
    w.add_name(f"qwen3-tts-random-{preset}")
    w.add_uint32(f"{arch}.block_count", p["n_layers"])
    w.add_uint32(f"{arch}.embedding_length", H)
    w.add_uint32(f"{arch}.feed_forward_length", FFN)
    w.add_uint32(f"{arch}.attention.head_count", NH)
    w.add_uint32(f"{arch}.attention.head_count_kv", NKV)
    w.add_uint32(f"{arch}.attention.key_length", HD)
    w.add_float32(f"{arch}.attention.layer_norm_rms_epsilon", 1e-6)
    w.add_float32(f"{arch}.rope.freq_base", 1000000.0)
    w.add_uint32(f"{arch}.vocab_size", CODEC_VOCAB)
    w.add_uint32(f"{arch}.text_vocab_size", p["text_vocab"])
    w.add_uint32(f"{arch}.text_hidden_size", TEXT_DIM)
    w.add_uint32(f"{arch}.num_code_groups", N_CB)
    w.add_uint32(f"{arch}.code_predictor.layer_count", p["cp_layers"])
    w.add_uint32(f"{arch}.code_predictor.vocab_size", CP_VOCAB)
    w.add_uint32(f"{arch}.codec.pad_id", 2148)
    w.add_uint32(f"{arch}.codec.bos_id", 2149)
    w.add_uint32(f"{arch}.codec.eos_id", 2150)
    w.add_uint32(f"{arch}.tts_bos_token_id", p["tts_bos"])
    w.add_uint32(f"{arch}.tts_eos_token_id", p["tts_eos"])
    w.add_uint32(f"{arch}.tts_pad_token_id", p["tts_pad"])

    if p["text_vocab"] >= 151646:
        # synthetic byte-level vocabulary for the reference's TextTokenizer (src/text_tokenizer.cpp:80-165), so the
        # unmodified qwen3-tts-cli can run on this file: the 256 GPT-2 byte tokens at ids 0..255 in
        # GPT-2 order ('!' = 0, 'Ġ' = 220, 'Ċ' = 198), "assistant" = 77091, <|im_start|> = 151644, <|im_end|> = 151645,
        # no merges -> text encodes byte by byte ("Hello, world!" -> HELLO_FULL of tests/util.py)
        w.add_token_list(synthetic_vocab(p["text_vocab"]))
        w.add_uint32("tokenizer.ggml.bos_token_id", 151644)
        w.add_uint32("tokenizer.ggml.eos_token_id", 151645)
        w.add_uint32("tokenizer.ggml.padding_token_id", 151643)

    q = wtype                                   # quantisable class (attn q/k/v/o, ffn gate/up/down, text_proj fc1/fc2)
    f = "f32" if wtype == "f32" else "f16"      # the rest: F16, or F32 in `--type f32` files (scripts/convert_tts_to_gguf.py:289-291)

    _add(w, "talker.text_embd.weight", rng.normal((p["text_vocab"], TEXT_DIM), w_std), f)
    _add(w, "talker.text_proj.fc1.weight", rng.normal((TEXT_DIM, TEXT_DIM), w_std), q)
    _add(w, "talker.text_proj.fc1.bias", rng.normal((TEXT_DIM,), 0.01), f)
    _add(w, "talker.text_proj.fc2.weight", rng.normal((H, TEXT_DIM), w_std), q)
    _add(w, "talker.text_proj.fc2.bias", rng.normal((H,), 0.01), f)
    _add(w, "talker.codec_embd.weight", rng.normal((CODEC_VOCAB, H), w_std), f)
    head = rng.normal((CODEC_VOCAB, H), w_std)
    if eos_like >= 0:
        head[2150] = head[eos_like] * np.float32(1.02)
    _add(w, "talker.codec_head.weight", head, f)
    _add(w, "talker.output_norm.weight", 1.0 + rng.normal((H,), 0.02), f)

    def layer(prefix):
        _add(w, f"{prefix}.attn_norm.weight", 1.0 + rng.normal((H,), 0.02), f)
        _add(w, f"{prefix}.attn_q.weight", rng.normal((NH * HD, H), w_std), q)
        _add(w, f"{prefix}.attn_k.weight", rng.normal((NKV * HD, H), w_std), q)
        _add(w, f"{prefix}.attn_v.weight", rng.normal((NKV * HD, H), w_std), q)
        _add(w, f"{prefix}.attn_output.weight", rng.normal((H, NH * HD), w_std), q)
        _add(w, f"{prefix}.attn_q_norm.weight", 1.0 + rng.normal((HD,), 0.02), f)
        _add(w, f"{prefix}.attn_k_norm.weight", 1.0 + rng.normal((HD,), 0.02), f)
        _add(w, f"{prefix}.ffn_norm.weight", 1.0 + rng.normal((H,), 0.02), f)
        _add(w, f"{prefix}.ffn_gate.weight", rng.normal((FFN, H), w_std), q)
        _add(w, f"{prefix}.ffn_up.weight", rng.normal((FFN, H), w_std), q)
        _add(w, f"{prefix}.ffn_down.weight", rng.normal((H, FFN), w_std), q)

    for il in range(p["n_layers"]):
        layer(f"talker.blk.{il}")
    for il in range(p["cp_layers"]):
        layer(f"code_pred.blk.{il}")
    for cb in range(N_CB - 1):
        _add(w, f"code_pred.codec_embd.{cb}.weight", rng.normal((CP_VOCAB, H), w_std), f)
    for cb in range(N_CB - 1):
        _add(w, f"code_pred.lm_head.{cb}.weight", rng.normal((CP_VOCAB, H), w_std), f)
    _add(w, "code_pred.output_norm.weight", 1.0 + rng.normal((H,), 0.02), f)

    add_speaker_encoder(w, seed)

    w.write_header_to_file(path=path)
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    return dict(preset=preset, seed=seed, wtype=wtype, **p)


def add_speaker_encoder(w, seed: int) -> None:
    """ECAPA-TDNN speaker encoder tensors (`spk_enc.*`, names and shapes of scripts/convert_tts_to_gguf.py:59-125 and
    src/audio_tokenizer_encoder.cpp:172-246; conv weights are numpy [Cout, Cin, K] -> F16, biases F32).  They come from
    a generator of their own (seed + 7919), after every other tensor, so the talker / code-predictor tensors keep their
    values and both presets carry the same encoder."""
    rng = _Rng(seed + 7919)
    w.add_uint32("qwen3-tts.speaker_encoder.embedding_length", 1024)
    w.add_uint32("qwen3-tts.speaker_encoder.sample_rate", 24000)

    def conv(name, cout, cin, k, gain=1.0):
        _add(w, f"spk_enc.{name}.weight", rng.normal((cout, cin, k), gain / np.sqrt(cin * k)), "f16")
        _add(w, f"spk_enc.{name}.bias", rng.normal((cout,), 0.05), "f16")

    conv("conv0", 512, 128, 5, gain=0.3)                 # log-mel inputs are O(5)
    for b in (1, 2, 3):
        conv(f"blk.{b}.tdnn1", 512, 512, 1, gain=1.4)
        for r in range(7):
            conv(f"blk.{b}.res2net.{r}", 64, 64, 3, gain=1.4)
        conv(f"blk.{b}.tdnn2", 512, 512, 1, gain=1.4)
        conv(f"blk.{b}.se.conv1", 128, 512, 1)
        conv(f"blk.{b}.se.conv2", 512, 128, 1)
    conv("mfa", 1536, 1536, 1, gain=1.4)
    conv("asp.tdnn", 128, 4608, 1)
    conv("asp.conv", 1536, 128, 1, gain=3.0)
    conv("fc", 1024, 3072, 1, gain=0.2)


def write_vocoder_gguf(path: str, seed: int = 4321, gain: float = 1.0, out_gain: float = 0.06, usage: bool = False,
                       wtype: str = "f16") -> dict:
    """Vocoder file (reference name: qwen3-tts-tokenizer-f16.gguf), `tok_dec.*` tensors only.

    Conv / linear weights ~ N(0, gain^2 / fan_in); snake alpha/beta ~ N(0, 0.1^2) (log-scale
    parameters, reference src/audio_tokenizer_decoder.cpp:383-397); layer scales 0.01-ish;
    codebooks N(0,1).  `out_gain` scales the final 96->1 conv so the pre-tanh RMS stays well
    inside the linear range of tanh (otherwise an SNR test is meaningless).

    usage=True appends the `*.usage` tensors of older converters (src/audio_tokenizer_decoder.cpp:40-73,141,189:
    codebook rows are divided by max(usage, 1e-5) at load); they come last so every other tensor keeps its values.
    """
    rng = _Rng(seed)
    arch = "qwen3-tts-tokenizer"
    w = gguf.GGUFWriter(path=None, arch=arch)
    w.add_name("qwen3-tts-tokenizer-random")
    w.add_uint32(f"{arch}.num_codebooks", 16)
    w.add_uint32(f"{arch}.codebook_size", 2048)
    w.add_uint32(f"{arch}.sample_rate", 24000)

    f = "f16"
    if wtype != "f16":
        # what the reference's converter does for --type q8_0 / f32 (scripts/convert_tokenizer_to_gguf.py:277-297): 1-D stay F32;
        # q8_0: codebook / norm / scale / alpha / beta names stay F16, everything else is handed to the Q8_0 quantiser, which
        # only succeeds when the last axis is a multiple of 32 (the conv kernels' K = 1..16 fall back to F16)
        real_add = _add

        def _add_v(writer, name, data, _q):
            if data.ndim <= 1:
                return real_add(writer, name, data, "f16")
            if wtype == "f32":
                return real_add(writer, name, data, "f32")
            keep = any(x in name for x in ["codebook", "_norm", "norm.", "scale", "alpha", "beta"])
            if keep or data.shape[-1] % 32 != 0:
                return real_add(writer, name, data, "f16")
            return real_add(writer, name, data, "q8_0")
        globals()["_add_voc"] = _add_v
    else:
        globals()["_add_voc"] = _add

    def fan(shape, fan_in, g=gain):
        return rng.normal(shape, g / np.sqrt(fan_in))

    _add_voc(w, "tok_dec.vq_first.input_proj.weight", fan((256, 512, 1), 512), f)
    _add_voc(w, "tok_dec.vq_first.output_proj.weight", fan((512, 256, 1), 256 * 16), f)
    def codebook():
        cb = rng.normal((2048, 256), 1.0)
        if usage:
            cb[::97] *= np.float32(1e-3)          # rows whose usage is below the loader's epsilon: x 1e5 must stay finite in F16
        return cb

    _add_voc(w, "tok_dec.vq_first.0.codebook", codebook(), f)
    _add_voc(w, "tok_dec.vq_rest.input_proj.weight", fan((256, 512, 1), 512), f)
    _add_voc(w, "tok_dec.vq_rest.output_proj.weight", fan((512, 256, 1), 256 * 16), f)
    for i in range(15):
        _add_voc(w, f"tok_dec.vq_rest.{i}.codebook", codebook(), f)
    _add_voc(w, "tok_dec.pre_conv.weight", fan((1024, 512, 3), 512 * 3), f)
    _add_voc(w, "tok_dec.pre_conv.bias", rng.normal((1024,), 0.01), f)
    _add_voc(w, "tok_dec.pre_tfm.input_proj.weight", fan((512, 1024), 1024), f)
    _add_voc(w, "tok_dec.pre_tfm.input_proj.bias", rng.normal((512,), 0.01), f)
    for i in range(8):
        b = f"tok_dec.pre_tfm.blk.{i}"
        _add_voc(w, f"{b}.attn_norm.weight", 1.0 + rng.normal((512,), 0.02), f)
        _add_voc(w, f"{b}.attn_q.weight", fan((1024, 512), 512), f)
        _add_voc(w, f"{b}.attn_k.weight", fan((1024, 512), 512), f)
        _add_voc(w, f"{b}.attn_v.weight", fan((1024, 512), 512), f)
        _add_voc(w, f"{b}.attn_output.weight", fan((512, 1024), 1024), f)
        _add_voc(w, f"{b}.attn_scale", 0.01 + rng.normal((512,), 0.002), f)
        _add_voc(w, f"{b}.ffn_norm.weight", 1.0 + rng.normal((512,), 0.02), f)
        _add_voc(w, f"{b}.ffn_gate.weight", fan((1024, 512), 512), f)
        _add_voc(w, f"{b}.ffn_up.weight", fan((1024, 512), 512), f)
        _add_voc(w, f"{b}.ffn_down.weight", fan((512, 1024), 1024), f)
        _add_voc(w, f"{b}.ffn_scale", 0.01 + rng.normal((512,), 0.002), f)
    _add_voc(w, "tok_dec.pre_tfm.norm.weight", 1.0 + rng.normal((512,), 0.02), f)
    _add_voc(w, "tok_dec.pre_tfm.output_proj.weight", fan((1024, 512), 512), f)
    _add_voc(w, "tok_dec.pre_tfm.output_proj.bias", rng.normal((1024,), 0.01), f)
    for i in range(2):
        b = f"tok_dec.upsample.{i}"
        _add_voc(w, f"{b}.conv.weight", fan((1024, 1024, 2), 1024), f)      # ConvT [Cin,Cout,K]
        _add_voc(w, f"{b}.conv.bias", rng.normal((1024,), 0.01), f)
        _add_voc(w, f"{b}.dwconv.weight", fan((1024, 1, 7), 7), f)
        _add_voc(w, f"{b}.dwconv.bias", rng.normal((1024,), 0.01), f)
        _add_voc(w, f"{b}.norm.weight", 1.0 + rng.normal((1024,), 0.02), f)
        _add_voc(w, f"{b}.norm.bias", rng.normal((1024,), 0.01), f)
        _add_voc(w, f"{b}.pwconv1.weight", fan((4096, 1024), 1024), f)
        _add_voc(w, f"{b}.pwconv1.bias", rng.normal((4096,), 0.01), f)
        _add_voc(w, f"{b}.pwconv2.weight", fan((1024, 4096), 4096), f)
        _add_voc(w, f"{b}.pwconv2.bias", rng.normal((1024,), 0.01), f)
        _add_voc(w, f"{b}.gamma", 0.1 + rng.normal((1024,), 0.01), f)
    _add_voc(w, "tok_dec.dec.0.conv.weight", fan((1536, 1024, 7), 1024 * 7), f)
    _add_voc(w, "tok_dec.dec.0.conv.bias", rng.normal((1536,), 0.01), f)
    cin = 1536
    for bi, s in enumerate((8, 5, 4, 3), start=1):
        cout = cin // 2
        b = f"tok_dec.dec.{bi}"
        _add_voc(w, f"{b}.snake.alpha", rng.normal((cin,), 0.1), f)
        _add_voc(w, f"{b}.snake.beta", rng.normal((cin,), 0.1), f)
        # each output sample sees K/s = 2 taps x Cin inputs
        _add_voc(w, f"{b}.conv_t.weight", fan((cin, cout, 2 * s), cin * 2), f)
        _add_voc(w, f"{b}.conv_t.bias", rng.normal((cout,), 0.01), f)
        for r in (2, 3, 4):
            rb = f"{b}.res.{r}"
            _add_voc(w, f"{rb}.act1.alpha", rng.normal((cout,), 0.1), f)
            _add_voc(w, f"{rb}.act1.beta", rng.normal((cout,), 0.1), f)
            _add_voc(w, f"{rb}.conv1.weight", fan((cout, cout, 7), cout * 7), f)
            _add_voc(w, f"{rb}.conv1.bias", rng.normal((cout,), 0.01), f)
            _add_voc(w, f"{rb}.act2.alpha", rng.normal((cout,), 0.1), f)
            _add_voc(w, f"{rb}.act2.beta", rng.normal((cout,), 0.1), f)
            _add_voc(w, f"{rb}.conv2.weight", fan((cout, cout, 1), cout, g=0.5 * gain), f)
            _add_voc(w, f"{rb}.conv2.bias", rng.normal((cout,), 0.01), f)
        cin = cout
    _add_voc(w, "tok_dec.dec.5.snake.alpha", rng.normal((96,), 0.1), f)
    _add_voc(w, "tok_dec.dec.5.snake.beta", rng.normal((96,), 0.1), f)
    _add_voc(w, "tok_dec.dec.6.conv.weight", fan((1, 96, 7), 96 * 7, g=out_gain), f)
    _add_voc(w, "tok_dec.dec.6.conv.bias", rng.normal((1,), 0.01), f)
    if usage:
        for name in ["tok_dec.vq_first.0.usage"] + [f"tok_dec.vq_rest.{i}.usage" for i in range(15)]:
            u = np.abs(rng.normal((2048,), 0.5)) + np.float32(0.5)
            u[::97] = np.float32(1e-7)                      # below the epsilon: clamped to 1e-5 by the loader
            _add_voc(w, name, u, f)

    w.write_header_to_file(path=path)
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    return dict(seed=seed, gain=gain, out_gain=out_gain)


def ensure_models(model_dir: str, preset: str = "tiny", wtype: str = "f16", seed: int = 1234,
                  voc_seed: int = 4321, verbose: bool = False, eos_like: int = -1,
                  usage: bool = False, voc_wtype: str = "f16") -> tuple[str, str]:
    """Create (if missing) the two GGUFs under model_dir; returns (tts_path, vocoder_path).

    File names follow the reference pipeline (src/qwen3_tts.cpp:118-119) with a preset /
    weight-type infix so several variants can sit side by side in one cache directory.
    """
    os.makedirs(model_dir, exist_ok=True)
    eos = "" if eos_like < 0 else f"-eoslike{eos_like}"
    tts = os.path.join(model_dir, f"qwen3-tts-0.6b-{preset}-{wtype}-s{seed}{eos}.gguf")
    voc = os.path.join(model_dir, f"qwen3-tts-tokenizer-{voc_wtype}-s{voc_seed}{'-usage' if usage else ''}.gguf")
    if not os.path.exists(tts):
        if verbose:
            print(f"[make_random_gguf] writing {tts}", file=sys.stderr)
        write_tts_gguf(tts + ".tmp", preset=preset, seed=seed, wtype=wtype, eos_like=eos_like)
        os.replace(tts + ".tmp", tts)
    if not os.path.exists(voc):
        if verbose:
            print(f"[make_random_gguf] writing {voc}", file=sys.stderr)
        write_vocoder_gguf(voc + ".tmp", seed=voc_seed, usage=usage, wtype=voc_wtype)
        os.replace(voc + ".tmp", voc)
    return tts, voc


def main():
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--dir", required=True)
    ap.add_argument("--preset", default="tiny", choices=list(PRESETS))
    ap.add_argument("--wtype", default="f16", choices=["f16", "q8_0", "q4_k", "f32"])
    ap.add_argument("--seed", type=int, default=1234)
    a = ap.parse_args()
    print(*ensure_models(a.dir, a.preset, a.wtype, a.seed, verbose=True))


if __name__ == "__main__":
    main()
