"""Request-parallel dispatch of independent utterances over the GPUs of one box.

The hot path shards by *utterance*: every rank (one process per GPU) holds a full replica of the
weights and synthesises the utterances it owns; there is no data-path collective (nothing to
reduce).  `torch.distributed` is used only to agree on the work list and to collect per-rank
results / timings on rank 0.  The reference API is one utterance at a time
(src/qwen3_tts.h:94-112); this is the front door for config 4 (64 concurrent utterances).

The engine is injected (`synth(tokens, spk, max_len) -> (codes, wav)`) so the partition / gather
logic is testable on CPU with the gloo backend and a stand-in engine.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

import numpy as np


@dataclass
class Utterance:
    uid: int
    tokens: np.ndarray                 # int32 text-token ids (>= 4)
    spk: Optional[np.ndarray] = None   # float32[1024] speaker embedding or None
    max_len: int = 320


def owner(uid_index: int, world: int) -> int:
    """Static partition of the benchmark: utterance i -> rank i mod G."""
    return uid_index % world


def my_share(utts: Sequence[Utterance], rank: int, world: int) -> List[int]:
    return [i for i in range(len(utts)) if owner(i, world) == rank]


def make_workload(n_utt: int, seed: int = 4, max_len: int = 320) -> List[Utterance]:
    """Config 4: texts of 8..64 bytes drawn from a fixed list with a seed, byte-level token ids in
    the synthetic vocabulary, wrapped in the chat template of the reference CLI."""
    rng = np.random.default_rng(seed)
    base = [b"Hello, world!", b"The quick brown fox jumps over the lazy dog.", b"Speech synthesis on one GPU.",
            b"How are you today?", b"Twelve and a half frames per second.", b"Request parallel, no collective.",
            b"A longer sentence gives the talker more trailing text rows to add.", b"OK."]
    out = []
    for i in range(n_utt):
        txt = base[int(rng.integers(0, len(base)))]
        n = int(rng.integers(8, 65))
        body = (txt * (n // len(txt) + 1))[:n]
        toks = [151644, 77091, 198] + [int(b) for b in body] + [151645, 198, 151644, 77091, 198]
        spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
        out.append(Utterance(i, np.asarray(toks, np.int32), spk, max_len))
    return out


def run_sharded(utts: Sequence[Utterance], synth: Optional[Callable], rank: int = 0, world: int = 1, dist=None,
                synth_batch: Optional[Callable] = None):
    """Synthesise this rank's share; gather {uid: (n_frames, n_samples, checksum)} on rank 0.

    `synth(tokens, spk, max_len) -> (codes, wav)` handles one utterance at a time;
    `synth_batch([Utterance, ...]) -> [(codes, wav), ...]` (preferred when given) hands the rank's whole
    share to the engine at once so it can batch the sequences on the device (lm_batch_kernel).
    Returns (local_results, gathered) where gathered is None on ranks != 0.  The per-utterance
    checksum (sum of codes) lets rank 0 verify that every utterance was produced exactly once.
    """
    local = {}
    mine = [utts[i] for i in my_share(utts, rank, world)]
    if synth_batch is not None and mine:
        outs = synth_batch(mine)
        if len(outs) != len(mine):
            raise RuntimeError("synth_batch returned a different number of results")
    else:
        outs = [synth(u.tokens, u.spk, u.max_len) for u in mine]
    for u, (codes, wav) in zip(mine, outs):
        local[u.uid] = (int(codes.shape[0]), int(len(wav)), int(np.asarray(codes, np.int64).sum()))
    if dist is None or world == 1:
        return local, dict(local)
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(local, gathered, dst=0)
    if rank != 0:
        return local, None
    merged = {}
    for part in gathered:
        for k, v in part.items():
            if k in merged:
                raise RuntimeError(f"utterance {k} produced by two ranks")
            merged[k] = v
    missing = [u.uid for u in utts if u.uid not in merged]
    if missing:
        raise RuntimeError(f"utterances not produced: {missing}")
    return local, merged
