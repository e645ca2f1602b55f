"""Request-parallel dispatcher (qwen3-tts.cpp_b200/dispatch.py) — host logic, CPU only.

world_size 2 over the gloo backend with a stand-in engine: the partition is i mod G, every
utterance is produced exactly once, rank 0 sees all results, and nothing but the work list and
the result summaries crosses ranks (no data-path collective)."""
import os
import socket
import sys

import numpy as np
import pytest

from util import PKG_DIR, pkg


def _load_dispatch():
    pkg()
    import importlib
    return importlib.import_module("q3t_b200.dispatch")


def _fake_synth(tokens, spk, max_len):
    # deterministic stand-in: the "codes" depend only on the utterance
    n = 3 + int(tokens.sum()) % 5
    codes = (np.arange(n * 16, dtype=np.int32).reshape(n, 16) + int(tokens[3])) % 2048
    return codes, np.zeros(1920 * n - 555, np.float32)


def test_partition_is_static_round_robin():
    d = _load_dispatch()
    utts = d.make_workload(10, seed=1, max_len=4)
    assert [d.owner(i, 4) for i in range(10)] == [0, 1, 2, 3, 0, 1, 2, 3, 0, 1]
    shares = [d.my_share(utts, r, 4) for r in range(4)]
    assert sorted(sum(shares, [])) == list(range(10))
    for u in utts:
        assert u.tokens.dtype == np.int32 and len(u.tokens) >= 8 + 8 and u.tokens[0] == 151644
        assert u.spk.shape == (1024,)
    a = d.make_workload(10, seed=1, max_len=4)
    assert all(np.array_equal(x.tokens, y.tokens) for x, y in zip(a, utts))      # seeded


def test_single_rank_runs_everything():
    d = _load_dispatch()
    utts = d.make_workload(7, seed=2, max_len=4)
    local, merged = d.run_sharded(utts, _fake_synth)
    assert sorted(merged) == list(range(7)) and merged == local


def test_batched_hand_over_gives_the_same_summaries():
    """synth_batch receives the rank's whole share at once (the engine batches it on the device)."""
    d = _load_dispatch()
    utts = d.make_workload(7, seed=2, max_len=4)
    seen = []

    def fake_batch(mine):
        seen.append([u.uid for u in mine])
        return [_fake_synth(u.tokens, u.spk, u.max_len) for u in mine]

    _, one = d.run_sharded(utts, _fake_synth)
    _, bat = d.run_sharded(utts, None, synth_batch=fake_batch)
    assert bat == one and seen == [list(range(7))]
    _, part = d.run_sharded(utts, None, 1, 2, None, synth_batch=fake_batch)        # rank 1 of 2, no gather
    assert seen[-1] == [1, 3, 5] and sorted(part) == [1, 3, 5]
    with pytest.raises(RuntimeError):
        d.run_sharded(utts, None, synth_batch=lambda mine: [])


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    d = _load_dispatch()
    utts = d.make_workload(9, seed=3, max_len=4)
    local, merged = d.run_sharded(utts, _fake_synth, rank, world, dist)
    q.put((rank, sorted(local), None if merged is None else sorted(merged.items())))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_gloo():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = {}
    for _ in range(2):
        r, local, merged = q.get(timeout=120)
        res[r] = (local, merged)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][0] == [0, 2, 4, 6, 8] and res[1][0] == [1, 3, 5, 7]
    assert res[1][1] is None
    d = _load_dispatch()
    utts = d.make_workload(9, seed=3, max_len=4)
    want = sorted((u.uid, (lambda c, w: (c.shape[0], len(w), int(c.astype(np.int64).sum())))(*_fake_synth(u.tokens, u.spk, 4)))
                  for u in utts)
    assert res[0][1] == want
