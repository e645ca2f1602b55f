"""q3t_project_text_static_dev — the allocation-free variant of the text projection (scratch in a
file-scope `static __device__` array of the library) — against the existing q3t_project_text entry
point, through the C ABI.  B200 only.  Collected last on purpose
(`zz`).

The two entry points launch the same mat-vec kernels on the same weights, so the comparison is bit
equality (the tolerance of the other parity tests of this path).  Two calls in a row with different
inputs check that the shared scratch array, which keeps the previous call's contents, does not leak
into the next result.  Device memory for the inputs and outputs is the CALLER's (torch here); the
entry point itself allocates nothing.
"""
import ctypes as C

import numpy as np
import pytest

from util import models

pytestmark = pytest.mark.gpu


def _project_host(q3t, t, ids):
    out = np.zeros((len(ids), t.cfg.hidden_size), np.float32)
    rc = q3t.lib().q3t_project_text(t.engine(), ids.ctypes.data_as(C.POINTER(C.c_int32)), len(ids),
                                    out.ctypes.data_as(C.POINTER(C.c_float)))
    assert rc == 0, q3t.lib().q3t_last_error(t.engine()).decode()
    return out


def _project_static(q3t, t, ids_dev, out_dev):
    return q3t.lib().q3t_project_text_static_dev(t.engine(), C.c_void_p(ids_dev.data_ptr()), int(ids_dev.numel()),
                                                 C.c_void_p(out_dev.data_ptr()))


@pytest.mark.parametrize("wtype", ["f16", "q8_0"])
def test_static_scratch_projection_matches_allocating_entry_point(q3t, wtype):
    import torch
    t = q3t.TTSTransformer()
    assert t.load_model(models("tiny", wtype)[0]), t.get_error()
    try:
        L = q3t.lib()
        vocab, Hd = t.cfg.text_vocab_size, t.cfg.hidden_size
        max_tok = L.q3t_project_text_static_max_tokens()
        assert max_tok == 64
        rng = np.random.default_rng(11)
        dev = torch.device("cuda:0")
        out_dev = torch.empty((max_tok, Hd), dtype=torch.float32, device=dev)
        # two calls in a row: a full-size one, then a short one with other ids (stale scratch rows behind it)
        for n in (max_tok, 7):
            ids = rng.integers(0, vocab, size=n).astype(np.int32)
            want = _project_host(q3t, t, ids)
            ids_dev = torch.from_numpy(ids).to(dev)
            out_dev.fill_(float("nan"))
            torch.cuda.synchronize()
            rc = _project_static(q3t, t, ids_dev, out_dev)
            assert rc == 0, L.q3t_last_error(t.engine()).decode()
            got = out_dev[:n].cpu().numpy()
            assert np.isfinite(got).all()
            assert np.array_equal(got, want), f"n={n}: max abs diff {np.abs(got - want).max()}"
            assert torch.isnan(out_dev[n:]).all()                 # rows beyond n_tokens are not touched
        # sizes beyond the static scratch are rejected, with a message
        big = torch.zeros(max_tok + 1, dtype=torch.int32, device=dev)
        big_out = torch.empty((max_tok + 1, Hd), dtype=torch.float32, device=dev)
        assert _project_static(q3t, t, big, big_out) == -1        # Q3T_ERR_ARG
        assert "exceeds the static scratch" in L.q3t_last_error(t.engine()).decode()
        # an id outside the vocabulary is caught on the device
        bad = torch.tensor([1, vocab, 2], dtype=torch.int32, device=dev)
        assert _project_static(q3t, t, bad, out_dev) == -1
        assert "out of range" in L.q3t_last_error(t.engine()).decode()
        # ... and the flag does not stick: the next good call succeeds and is exact again
        ids = rng.integers(0, vocab, size=5).astype(np.int32)
        assert _project_static(q3t, t, torch.from_numpy(ids).to(dev), out_dev) == 0
        assert np.array_equal(out_dev[:5].cpu().numpy(), _project_host(q3t, t, ids))
    finally:
        t.close()
