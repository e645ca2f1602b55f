"""Shared helpers for the test-suite: package import (the directory name is not a Python
identifier), model cache, reference access."""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_DIR = os.path.join(ROOT, "qwen3-tts.cpp_b200")
MODEL_DIR = os.environ.get("Q3T_MODEL_DIR", "/tmp/q3t_models")
GOLDEN = os.path.join(ROOT, "tests", "golden")

if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pkg():
    if "q3t_b200" in sys.modules:
        return sys.modules["q3t_b200"]
    spec = importlib.util.spec_from_file_location("q3t_b200", os.path.join(PKG_DIR, "__init__.py"),
                                                  submodule_search_locations=[PKG_DIR])
    m = importlib.util.module_from_spec(spec)
    sys.modules["q3t_b200"] = m
    spec.loader.exec_module(m)
    return m


def models(preset="tiny", wtype="f16", **kw):
    from tools.make_random_gguf import ensure_models
    return ensure_models(MODEL_DIR, preset=preset, wtype=wtype, **kw)


# "Hello, world!" with the synthetic byte-level vocabulary, remapped into
# the tiny preset's 512-entry text vocabulary where needed.
HELLO_FULL = np.array([151644, 77091, 198, 39, 68, 75, 75, 78, 11, 220, 86, 78, 81, 75, 67, 0,
                       151645, 198, 151644, 77091, 198], np.int32)
HELLO_TINY = np.array([300, 301, 198, 39, 68, 75, 75, 78, 11, 220, 86, 78, 81, 75, 67, 0,
                       302, 198, 300, 301, 198], np.int32)


def snr_db(ref, x):
    ref = np.asarray(ref, np.float64); x = np.asarray(x, np.float64)
    n = min(len(ref), len(x))
    err = ref[:n] - x[:n]
    return 10.0 * np.log10((ref[:n] ** 2).sum() / max((err ** 2).sum(), 1e-300))
