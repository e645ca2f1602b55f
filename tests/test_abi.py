"""C-ABI boundary (include/q3t.h) and host logic — runs without a GPU.

The library must load, export every symbol the header declares, and every compute entry point
must fail loudly (no CPU fallback) when there is no CUDA device.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from util import ROOT


def test_library_exports_every_declared_symbol(q3t):
    L = q3t.lib()
    names = q3t.declared_symbols()
    assert len(names) >= 60
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, f"declared in include/q3t.h but not exported: {missing}"


def test_header_is_plain_c():
    """include/q3t.h must compile as C (no C++/torch types at the boundary)."""
    src = '#include "q3t.h"\nint main(void){ return sizeof(q3t_tts_config) > 0 ? 0 : 1; }\n'
    r = subprocess.run(["gcc", "-std=c99", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), "-x", "c", "-"],
                       input=src, text=True, capture_output=True)
    assert r.returncode == 0, r.stderr


def test_no_oracle_reference_in_product():
    """The product tree never imports / links anything under oracle/."""
    pkg_dir = os.path.join(ROOT, "qwen3-tts.cpp_b200")
    for dp, _, fs in os.walk(pkg_dir):
        if "build" in dp:
            continue
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "libq3t_oracle" not in txt and "libq3t_ref" not in txt and "import oracle" not in txt \
                    and "from oracle" not in txt, f"{f} refers to the oracle"


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_has_cuda(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_gpu(q3t, tiny_models):
    L = q3t.lib()
    ctx = C.c_void_p()
    L.q3t_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    rc = L.q3t_ctx_create(0, C.byref(ctx))
    assert rc < 0 and not ctx.value
    assert b"no CPU fallback" in L.q3t_last_error(None)
    t = q3t.TTSTransformer()
    assert t.load_model(tiny_models[0]) is False
    assert "Failed to initialize backend" in t.get_error()
    d = q3t.AudioTokenizerDecoder()
    assert d.load_model(tiny_models[1]) is False
    assert d.decode(np.zeros((2, 16), np.int32)) is None
    assert d.get_error() == "Model not loaded"
    assert t.generate(np.arange(8, dtype=np.int32), None, 4) is None
    assert t.get_error() == "Model not loaded"
    # the extensions fail the same way (no model, no device): streaming decode, legacy predict_codes, batching
    d.stream_reset()
    assert d.decode_append(np.zeros((2, 16), np.int32)) is None
    assert d.get_error() == "Model not loaded"
    assert t.predict_codes(np.zeros(1024, np.float32), None) is None
    assert t.get_error() == "Model not loaded"
    assert t.generate_batch([(np.arange(8, dtype=np.int32), None)], 4) is None
    # the batched decode front door: an unloaded decoder is reported, bad arguments are rejected before any decoder is touched
    with pytest.raises(q3t.Q3TError, match="Model not loaded"):
        q3t.decode_batch([d], [np.zeros((2, 16), np.int32)])
    assert L.q3t_pipe_decode_batch_multi(None, 0, 0, None, None, None, None, None) == 0
    assert L.q3t_pipe_get_error() == b"bad argument"


def test_gguf_loader_matches_python_reader(q3t, tiny_models):
    """Own GGUF v3 reader (host/b200_gguf_reader.cpp) vs the `gguf` python package on the same file."""
    import gguf
    for path in tiny_models:
        g = q3t.GGUFLoader()
        assert g.open(path), g.get_error()
        r = gguf.GGUFReader(path)
        assert g.get_n_tensors() == len(r.tensors)
        assert g.get_data_offset() == r.data_offset
        for i, t in enumerate(r.tensors):
            assert g.get_tensor_name(i) == t.name
            assert g.get_tensor_type(i) == int(t.tensor_type)
            assert g.get_tensor_size(i) == int(t.n_bytes)
            assert g.get_tensor_offset(i) + r.data_offset == int(t.data_offset)
            ne = g.get_tensor_ne(i)
            shape = [int(x) for x in t.shape]
            assert ne[:len(shape)] == shape and all(x == 1 for x in ne[len(shape):])
        g.close()
    g = q3t.GGUFLoader()
    assert g.open(tiny_models[0])
    assert g.get_u32("qwen3-tts.block_count", -1) == 2
    assert g.get_u32("qwen3-tts.no_such_key", 77) == 77
    assert abs(g.get_f32("qwen3-tts.rope.freq_base", 0.0) - 1e6) < 1e-3
    assert abs(g.get_f32("qwen3-tts.attention.layer_norm_rms_epsilon", 0.0) - 1e-6) < 1e-12


def test_gguf_loader_rejects_garbage(q3t, tmp_path):
    bad = tmp_path / "bad.gguf"
    bad.write_bytes(b"NOPE" + b"\0" * 64)
    g = q3t.GGUFLoader()
    assert g.open(str(bad)) is False and g.get_error()
    assert g.open(str(tmp_path / "missing.gguf")) is False
    trunc = tmp_path / "trunc.gguf"
    trunc.write_bytes(b"GGUF\x03\x00\x00\x00" + b"\xff" * 8)
    assert g.open(str(trunc)) is False


def _gguf_bytes(tensors, kvs=(), pad_data=64, alignment=None):
    """Hand-assembled GGUF v3 file: tensors = [(name, ne, type, offset)], kvs = [(key, type, payload bytes)]."""
    import struct
    def s(x): return struct.pack("<Q", len(x)) + x.encode()
    kvs = list(kvs)
    if alignment is not None:
        kvs.append(("general.alignment", 4, struct.pack("<I", alignment)))
    out = b"GGUF" + struct.pack("<IQQ", 3, len(tensors), len(kvs))
    for key, typ, payload in kvs:
        out += s(key) + struct.pack("<I", typ) + payload
    for name, ne, typ, off in tensors:
        out += s(name) + struct.pack("<I", len(ne)) + b"".join(struct.pack("<Q", n & (2**64 - 1)) for n in ne)
        out += struct.pack("<IQ", typ, off & (2**64 - 1))
    a = alignment if alignment else 32
    out += b"\0" * ((-len(out)) % a)
    return out + b"\x11" * pad_data


def test_gguf_loader_rejects_overflowing_headers(q3t, tmp_path):
    """bounds checks must not be defeatable by 64-bit wrap-around (the reference's ggml parser rejects
    such files, ggml/src/gguf.cpp gguf_init_from_file_impl); every file here is refused at open(), none crashes."""
    import struct
    g = q3t.GGUFLoader()
    ok = tmp_path / "ok.gguf"
    ok.write_bytes(_gguf_bytes([("a", [8, 2], 0, 0)]))
    assert g.open(str(ok)), g.get_error()
    assert g.get_n_tensors() == 1 and g.get_tensor_size(0) == 64
    cases = {
        "offset_wraps":   _gguf_bytes([("a", [8, 2], 0, 2**64 - 32)]),            # data_offset + offset wraps to a small value
        "offset_past":    _gguf_bytes([("a", [8, 2], 0, 32)]),                     # 32 + 64 > 64 bytes of data
        "ne_product":     _gguf_bytes([("a", [2**32, 2**32], 0, 0)]),              # element count wraps to 0
        "ne_product3":    _gguf_bytes([("a", [2**21, 2**21, 2**21, 2], 1, 0)]),    # 2^64 halves
        "ne_negative":    _gguf_bytes([("a", [-8, 2], 0, 0)]),
        "ne_zero":        _gguf_bytes([("a", [0, 2], 0, 0)]),
        "blck_misfit":    _gguf_bytes([("a", [33, 2], 8, 0)]),                     # Q8_0 row not a multiple of 32
        "unknown_type":   _gguf_bytes([("a", [8, 2], 99, 0)]),
        "too_many_dims":  _gguf_bytes([("a", [2, 2, 2, 2, 2], 0, 0)]),
        "array_wraps":    _gguf_bytes([("a", [8, 2], 0, 0)], kvs=[("k", 9, struct.pack("<IQ", 10, 2**61 + 1))]),  # u64[] * 8 wraps
        "array_long":     _gguf_bytes([("a", [8, 2], 0, 0)], kvs=[("k", 9, struct.pack("<IQ", 4, 2**40))]),
        "string_long":    _gguf_bytes([("a", [8, 2], 0, 0)], kvs=[("k", 8, struct.pack("<Q", 2**63))]),
        "bad_kv_type":    _gguf_bytes([("a", [8, 2], 0, 0)], kvs=[("k", 77, b"\0" * 8)]),
        "alignment_zero": _gguf_bytes([("a", [8, 2], 0, 0)], alignment=0),
        "alignment_odd":  _gguf_bytes([("a", [8, 2], 0, 0)], alignment=48),
        "alignment_huge": _gguf_bytes([("a", [8, 2], 0, 0)], alignment=2**31),
        "counts_huge":    b"GGUF" + struct.pack("<IQQ", 3, 2**60, 2**60),
    }
    good = _gguf_bytes([("a", [8, 2], 0, 0)])
    for cut in (5, 12, 23, 24, 30, len(good) - 65, len(good) - 1):                 # truncations of a good file
        cases[f"trunc_{cut}"] = good[:cut]
    for name, blob in cases.items():
        f = tmp_path / f"{name}.gguf"
        f.write_bytes(blob)
        assert g.open(str(f)) is False, name
        assert g.get_error(), name
        assert g.get_n_tensors() == 0, name


def test_gguf_loader_alignment_is_per_file(q3t, tmp_path):
    """close()/re-open must not carry general.alignment from the previous file over."""
    g = q3t.GGUFLoader()
    a = tmp_path / "a64.gguf"; a.write_bytes(_gguf_bytes([("a", [8, 2], 0, 0)], alignment=256))
    b = tmp_path / "b.gguf";   b.write_bytes(_gguf_bytes([("a", [8, 2], 0, 0)]))
    assert g.open(str(a)), g.get_error()
    assert g.get_data_offset() % 256 == 0
    off_a = g.get_data_offset()
    assert g.open(str(b)), g.get_error()
    assert g.get_data_offset() % 32 == 0 and g.get_data_offset() < off_a
    big = tmp_path / "big.gguf"
    import struct
    big.write_bytes(_gguf_bytes([("a", [8, 2], 0, 0)], kvs=[("qwen3-tts.block_count", 10, struct.pack("<Q", 2**40 + 5)),
                                                              ("small", 10, struct.pack("<Q", 28))]))
    assert g.open(str(big)), g.get_error()
    assert g.get_u32("qwen3-tts.block_count", -7) == -7          # does not fit 32 bits: default, not 5
    assert g.get_u32("small", -7) == 28


def test_q8_0_model_metadata(q3t):
    from util import models
    tts, _ = models("tiny", "q8_0")
    g = q3t.GGUFLoader()
    assert g.open(tts)
    types = {g.get_tensor_name(i): g.get_tensor_type(i) for i in range(g.get_n_tensors())}
    assert types["talker.blk.0.attn_q.weight"] == 8          # Q8_0
    assert types["talker.codec_head.weight"] == 1            # stays F16
    assert types["talker.blk.0.attn_norm.weight"] == 0       # F32
    i = [g.get_tensor_name(k) for k in range(g.get_n_tensors())].index("talker.blk.0.attn_q.weight")
    assert g.get_tensor_size(i) == 2048 * (1024 // 32) * 34
