"""qwen3-tts.cpp_b200 — B200-native codes-to-audio hot path of qwen3-tts.cpp.

The product is the shared library ``libq3t_b200.so`` (CUDA engine + C++ host classes, C ABI in
``include/q3t.h``).  This package is only the Python view of that C ABI (ctypes), used by the
tests, ``bench.py`` and ``__graft_entry__``; it mirrors the reference's class names
(``TTSTransformer``, ``AudioTokenizerDecoder``, ``GGUFLoader``; src/tts_transformer.h:190-280,
src/audio_tokenizer_decoder.h:159-179, src/gguf_loader.h:16-62).

There is no CPU fallback anywhere: if the library is missing, import of ``binding`` raises, and
without a B200 every compute call returns the engine's error.
"""
from .binding import (  # noqa: F401
    LIB_PATH, lib, build, TTSTransformer, AudioTokenizerDecoder, GGUFLoader, Q3TError, declared_symbols,
    synthesize_stream, synthesize_batch, decode_batch, StreamStats, BatchStats, AudioTokenizerEncoder,
)
