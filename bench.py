#!/usr/bin/env python3
"""bench.py — codes-to-audio hot path of qwen3-tts.cpp on B200 (this repo) or on the host CPU
(the unmodified reference, --impl reference).

One *step* = one utterance through the hot path: prompt prefill, `--frames` greedy 12.5 Hz frames
(talker decode step + 15-pass code predictor each) and the vocoder over those frames.
Metric (BASELINE.json): audio-seconds generated per wall-second.

    python bench.py --gpus N --steps K --warmup W            # this repo, N ranks under torchrun for N>1
    python bench.py --impl reference --steps K --warmup W    # reference ggml-CPU path, bounded sample

Prints ONE JSON line on rank 0.  `value` = device-resident throughput (prompt already staged in
HBM, CUDA events on the engine stream); `e2e` = the same through TTSTransformer.generate +
AudioTokenizerDecoder.decode with host buffers; `roofline` = the persistent decode kernel
(one launch = one frame) against the measured HBM peak; `cpu_baseline` = oracle/_ref timed on
this box's host cores on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "audio_seconds_per_wall_second"
UNIT = "audio-s/s"
FRAME_RATE = 12.5

# algorithmic bytes (SURVEY.md §8d / DESIGN.md §5), F16 weights
TALKER_STEP_BYTES = {"f16": 887_357_440, "q8_0": 474_480_640}
CODEPRED_FRAME_BYTES = {"f16": 2_422_963_200, "q8_0": 1_317_043_200}
KV_BYTES_PER_POS = 114_688


def ncu_traffic(wtype: str):
    """dram bytes per lm_kernel launch from the committed ncu --set full capture (None if not captured for this dtype)."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "lm_kernel_traffic.json")))[wtype]
        return int(t["dram_bytes_read"]) + int(t["dram_bytes_write"])
    except Exception:
        return None


def n_samples(frames: int) -> int:
    return 1920 * frames - 555 if frames > 0 else 0


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index: int):
        self.index, self.rows, self.stop_ev, self.th = index, [], threading.Event(), None

    def _loop(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        while not self.stop_ev.is_set():
            try:
                r = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5)
                f = [x.strip() for x in r.stdout.strip().split(",")]
                if len(f) >= 6:
                    self.rows.append(f)
            except Exception:
                pass
            self.stop_ev.wait(0.2)

    def start(self):
        self.th = threading.Thread(target=self._loop, daemon=True)
        self.th.start()

    def stop(self):
        self.stop_ev.set()
        if self.th:
            self.th.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(int(float(r[0])) for r in self.rows)
        reasons = []
        for i, name in enumerate(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]):
            if any(r[2 + i].lower().startswith("active") for r in self.rows):
                reasons.append(name)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": int(float(self.rows[0][1])), "reasons": reasons,
                "samples": len(sm)}


def workload(frames):
    from util import HELLO_FULL
    spk_path = os.path.join(ROOT, "tests", "golden", "speaker_xvector.npz")
    if os.path.exists(spk_path):
        # config 2: the x-vector the reference's own ECAPA encoder gives for examples/readme_clone_input.wav with the
        # random-init spk_enc.* weights (golden G7, oracle/gen_golden.py --speaker)
        spk = np.load(spk_path)["embedding"].astype(np.float32)
    else:
        rng = np.random.default_rng(2024)
        spk = (rng.standard_normal(1024) * 0.05).astype(np.float32)
    return HELLO_FULL, spk, frames


def config_block(args, frames):
    """The `config` object: identical in both arms (same workload, same frames)."""
    decoding = "greedy" if args.temperature <= 0 else f"sampled (temp {args.temperature}, top-k {args.top_k}, rep-penalty 1.05)"
    return {"workload": f"0.6B {args.wtype.upper()} voice clone ('Hello, world!' + 1024-d speaker x-vector), {decoding}, {frames} frames "
                        f"(12.5 Hz, 16 codebooks) + vocoder over them = one utterance per step; this repo: batch 1 per GPU, one utterance "
                        f"per rank per step; reference arm: the same utterance on the host CPU (unmodified reference, ggml-CPU AVX2 build, "
                        f"all host threads, ONE CPU process regardless of --gpus)",
            "frames": frames, "samples": n_samples(frames), "batch_per_gpu": 1, "vocoder": True,
            "l2": "weights streamed per frame (3.3 GB) exceed the 126 MB L2; no explicit flush"}


# ------------------------------------------------------------------------------ reference arm
def run_reference(args, rank, world):
    if rank != 0:
        return
    from oracle import reflib
    from util import models
    if not reflib.available():
        emit_json(({"impl": "reference", "unavailable": "oracle/_ref/libq3t_ref.so was not built"}))
        return
    cores = os.cpu_count() or 1
    reflib.RefLib.get().set_threads(cores)
    tts_path, voc_path = models("full", args.wtype)
    tokens, spk, _ = workload(args.frames)
    sample_frames = args.ref_frames if args.ref_frames > 0 else args.frames
    ref = reflib.RefTTS(tts_path)
    dec = reflib.RefDecoder(voc_path)

    def step(n=sample_frames):
        codes = ref.generate(tokens, spk, n, temperature=args.temperature, top_k=args.top_k)
        wav = dec.decode(codes)
        return codes, wav

    # warm-up steps touch every weight page and spin the thread pool up; on the CPU that needs a few frames, not a
    # whole 25 s utterance (the timed steps below are the full workload)
    for _ in range(args.warmup):
        step(min(sample_frames, 8))
    t0 = time.perf_counter()
    for _ in range(args.steps):
        codes, wav = step()
    dt = time.perf_counter() - t0
    audio_s = len(wav) / 24000.0 * args.steps
    v = audio_s / dt
    sample = (f"{sample_frames} frames (generate) + vocoder over them per step = the whole workload of `config`, unmodified reference "
              f"(oracle/_ref, ggml-CPU, AVX2+FMA+F16C build, not -march=native), {cores} threads, one CPU process regardless of --gpus; "
              f"warm-up steps are 8-frame runs")
    emit_json(({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f16" if args.wtype == "f16" else "q8_0", "data": "synthetic",
        "config": config_block(args, sample_frames),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ------------------------------------------------------------------------------ this repo
def cpu_baseline(args):
    """oracle/_ref on the host cores, bounded sample (rank 0, N=1 only)."""
    try:
        from oracle import reflib
        from util import models
        if not reflib.available():
            return None
        cores = os.cpu_count() or 1
        reflib.RefLib.get().set_threads(cores)
        tts_path, voc_path = models("full", args.wtype)
        tokens, spk, _ = workload(args.frames)
        n = args.ref_frames if args.ref_frames > 0 else min(args.frames, 96)      # bounded sample (~10 s of CPU work)
        ref = reflib.RefTTS(tts_path)
        dec = reflib.RefDecoder(voc_path)
        ref.generate(tokens, spk, 2, temperature=0.0)                 # warm the page cache / threadpool
        t0 = time.perf_counter()
        codes = ref.generate(tokens, spk, n, temperature=0.0)
        t1 = time.perf_counter()
        wav = dec.decode(codes)
        t2 = time.perf_counter()
        ref.close(); dec.close()
        return {"value": (len(wav) / 24000.0) / (t2 - t0), "unit": UNIT, "cores": cores, "kind": "reference",
                "sample": f"{n} greedy frames + vocoder over them, ggml-CPU AVX2 build at {cores} threads "
                          f"(generate {(t1 - t0) / n * 1e3:.1f} ms/frame, vocoder {(t2 - t1) / n * 1e3:.1f} ms/frame)"}
    except Exception as e:                                            # the baseline never takes the bench down
        return {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "reference", "sample": f"failed: {e}"}


def run_ours(args, rank, world, local_rank):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this engine has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    os.environ["QWEN3_TTS_DEVICE"] = str(local_rank)

    from util import pkg, models
    m = pkg()
    if rank == 0:
        models("full", args.wtype)                                     # one rank writes the GGUFs
    if dist:
        dist.barrier()
    tts_path, voc_path = models("full", args.wtype)
    tokens, spk, frames = workload(args.frames)

    t = m.TTSTransformer()
    if not t.load_model(tts_path):
        raise SystemExit("load_model: " + t.get_error())
    d = m.AudioTokenizerDecoder()
    have_voc = (not args.no_vocoder) and d.load_model(voc_path)
    if not have_voc and not args.no_vocoder:
        raise SystemExit("decoder load_model: " + d.get_error())

    def sync_all():
        torch.cuda.synchronize()
        if dist:
            dist.barrier()
            torch.cuda.synchronize()

    if args.utterances > 0:
        return run_sharded_workload(args, rank, world, local_rank, t, d, have_voc, dist, sync_all, m)

    def step_resident():
        return t.run_staged(frames, repetition_penalty=1.05, temperature=args.temperature, top_k=args.top_k, seed=args.seed,
                            with_vocoder=have_voc)

    def step_e2e():
        codes = t.generate(tokens, spk, frames, repetition_penalty=1.05, temperature=args.temperature, top_k=args.top_k)
        if codes is None:
            raise SystemExit("generate: " + t.get_error())
        wav = d.decode(codes) if have_voc else np.zeros(n_samples(len(codes)), np.float32)
        if wav is None:
            raise SystemExit("decode: " + d.get_error())
        return codes, wav

    # ---- warm-up (both paths)
    t.stage_prompt(tokens, spk)
    for _ in range(max(args.warmup, 3)):
        step_resident()
    step_e2e()

    # ---- timed: device-resident (`value`)
    t.stage_prompt(tokens, spk)
    sampler = ClockSampler(local_rank)
    launches0 = t.timing().kernel_launches
    sync_all()
    sampler.start()
    w0 = time.perf_counter()
    dev_ms = prefill_ms = frames_ms = voc_ms = 0.0
    nf = ns = 0
    for _ in range(args.steps):
        nf, ns = step_resident()
        tm = t.timing()
        prefill_ms += tm.prefill_ms; frames_ms += tm.frames_ms; voc_ms += tm.vocoder_ms
    sync_all()
    wall_s = time.perf_counter() - w0
    clocks = sampler.stop()
    launches = t.timing().kernel_launches - launches0
    dev_ms = prefill_ms + frames_ms + voc_ms
    # the steps run back to back on one stream; the bracketed wall time is the honest whole-step figure
    step_s = wall_s / args.steps
    audio_s = (ns if have_voc else n_samples(nf)) / 24000.0

    # ---- timed: end to end through the host classes (`e2e`)
    sync_all()
    e0 = time.perf_counter()
    for _ in range(args.steps):
        codes, wav = step_e2e()
    sync_all()
    e2e_s = (time.perf_counter() - e0) / args.steps

    # ---- streaming front door: time from the call (host tokens) to the first audible samples on the host
    first_audio = None
    if have_voc:
        try:
            m.synthesize_stream(t, d, tokens, spk, frames, first_chunk=8, chunk=64)                   # warm
            fa = [m.synthesize_stream(t, d, tokens, spk, frames, first_chunk=8, chunk=64)[1] for _ in range(3)]
            first_audio = {"t_first_audio_ms": min(x.t_first_audio_ms for x in fa), "t_total_ms": min(x.t_total_ms for x in fa),
                           "first_chunk_frames": 8, "chunk_frames": 64, "frames": int(fa[0].n_frames),
                           "what": "q3t_pipe_synthesize_stream: prompt + prefill + 8 frames + streaming vocoder decode of those frames, "
                                   "host tokens in, host samples out (the reference decodes only after the last frame)"}
        except Exception as e:                                   # never takes the bench down
            first_audio = {"error": str(e)}

    # ---- speaker encoder (config 2's x-vector; the reference needs ~27 s of CPU for a 30 s clip, OPTIMIZATION.md:30)
    spk_enc = None
    if rank == 0:
        try:
            enc = m.AudioTokenizerEncoder()
            if enc.load_model(tts_path):
                rng = np.random.default_rng(7)
                clip = (0.1 * rng.standard_normal(24000 * 30)).astype(np.float32)
                enc.encode(clip)
                torch.cuda.synchronize(); s0 = time.perf_counter()
                enc.encode(clip)
                torch.cuda.synchronize()
                spk_enc = {"clip_seconds": 30, "ms": (time.perf_counter() - s0) * 1e3, "what": "host samples in, 1024-d x-vector out: log-mel "
                           "(DFT as an F32 GEMM with the reference's twiddles) + ECAPA-TDNN on the tap GEMM"}
            enc.close()
        except Exception as e:
            spk_enc = {"error": str(e)}

    # ---- talker decode step alone (north-star: HBM GB/s of the batch-1 talker step)
    n_mid = t.cfg_prefill_len if hasattr(t, "cfg_prefill_len") else 10
    us_step = t.bench_talker_step(n_mid + frames // 2, 64)

    times = torch.tensor([step_s, e2e_s], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    step_s, e2e_s = float(times[0]), float(times[1])

    # ---- BASELINE.json config 4 at this N: 64 utterances, strong scaling, batched engine (every rank takes part)
    c4 = None if args.no_batch else config4_line(args, rank, world, t, d, have_voc, dist, sync_all, m)

    if rank == 0:
        peak, peak_kind = measured_peaks()
        avg_kv = 10 + frames / 2.0
        frame_bytes = CODEPRED_FRAME_BYTES[args.wtype] + TALKER_STEP_BYTES[args.wtype] + KV_BYTES_PER_POS * (avg_kv + 1)
        frame_us = frames_ms / args.steps / max(nf, 1) * 1e3
        achieved = frame_bytes / (frame_us * 1e-6) / 1e9
        talker_bytes = TALKER_STEP_BYTES[args.wtype] + KV_BYTES_PER_POS * (n_mid + frames // 2 + 1)
        out = {
            "metric": METRIC, "value": world * audio_s / step_s, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f16" if args.wtype == "f16" else "q8_0 weights, int8 x int8 / f32 accumulate",
            "data": "synthetic (seeded random-init GGUF of the real 0.6B shapes; no checkpoint offline)",
            "config": config_block(args, frames),
            "parallelism": f"request-parallel x{world} (no collective)",
            "device_ms_per_step": {"prefill": prefill_ms / args.steps, "frames": frames_ms / args.steps,
                                   "vocoder": voc_ms / args.steps, "events_total": dev_ms / args.steps},
            "e2e": {"value": world * audio_s / e2e_s, "unit": UNIT,
                    "h2d_bytes_per_step": int(tokens.nbytes + spk.nbytes + codes.nbytes),
                    "d2h_bytes_per_step": int(codes.nbytes + wav.nbytes), "ms_per_step": e2e_s * 1e3},
            "gpu_launches": int(launches),
            "roofline": {"kernel": "lm_kernel (one launch = one frame: 16 code-predictor passes + 1 talker step)",
                         "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "peak_source": peak_kind, "traffic": ncu_traffic(args.wtype), "us_per_launch": frame_us,
                         "algorithmic_bytes_per_launch": int(frame_bytes)},
            "talker_step": {"us": us_step, "algorithmic_bytes": int(talker_bytes),
                            "achieved_gbs": talker_bytes / (us_step * 1e-6) / 1e9,
                            "frac": talker_bytes / (us_step * 1e-6) / 1e9 / peak, "n_kv": n_mid + frames // 2 + 1},
            "realtime_factor_per_gpu": audio_s / step_s,
            "first_audio": first_audio,
            "speaker_encoder": spk_enc,
            "clocks": clocks,
        }
        if c4 is not None:
            out["config4"] = c4
            if world == 1:
                out["batch64"] = c4                      # BASELINE.json's "batch 64" metric point = config 4 on one GPU
        if world == 1 and not args.no_cpu_baseline:
            cb = cpu_baseline(args)
            if cb:
                out["cpu_baseline"] = cb
        emit_json((out))
    if dist:
        dist.barrier()
        dist.destroy_process_group()


def config4_line(args, rank, world, t, d, have_voc, dist, sync_all, m):
    """BASELINE.json config 4 / the "batch 64" metric point: 64 utterances (8-64 byte texts, seed 4), greedy, `--frames`
    frames each + vocoder, strong scaling: utterance i -> rank i mod N, every rank pushes its share through the batched
    engine (generate_batch: 16 sequences per launch, two interleaved lock-step groups of 8), then decodes each utterance.  Host buffers in and out;
    one timed step bracketed by barrier + synchronize, max over ranks.  Returned on rank 0 (None elsewhere)."""
    import importlib
    import torch
    disp = importlib.import_module("q3t_b200.dispatch")
    n_utt = 64
    utts = disp.make_workload(n_utt, seed=4, max_len=args.frames)
    mine = [utts[i] for i in disp.my_share(utts, rank, world)]
    pairs = [(u.tokens, u.spk) for u in mine]
    if pairs and t.generate_batch(pairs[:min(8, len(pairs))], 8) is None:          # warm-up (instantiation, buffers)
        raise SystemExit("generate_batch: " + t.get_error())
    extra_dec = []
    if args.decoders <= 0:
        args.decoders = max(1, min(4, (os.cpu_count() or 8) // (2 * world)))
    if have_voc and args.decoders > 1:
        from util import models as _models
        for _ in range(args.decoders - 1):
            dd = m.AudioTokenizerDecoder()
            if not dd.load_model(_models("full", args.wtype)[1]):
                raise SystemExit("decoder load_model: " + dd.get_error())
            dd.decode(np.zeros((args.frames, 16), np.int32))                        # warm-up at the job's length (scratch sized: no cudaMalloc in the timed region)
            extra_dec.append(dd)
    sync_all()
    w0 = time.perf_counter()
    codes = t.generate_batch(pairs, args.frames, repetition_penalty=1.05) if pairs else []
    if codes is None:
        raise SystemExit("generate_batch: " + t.get_error())
    tm = t.timing()
    torch.cuda.synchronize()
    w1 = time.perf_counter()
    ns = 0
    if have_voc and len(extra_dec) > 0:
        # the utterances of the job are decoded by N decoder instances from N host threads (each instance has its own stream:
        # the small early-stage launches of one decode overlap the large late-stage GEMMs of another)
        ns = sum(len(w) for w in m.decode_batch([d] + extra_dec, codes))     # q3t_pipe_decode_batch_multi: the threads are C++ (no GIL)
    else:
        for c in codes:
            wav = d.decode(c) if have_voc else np.zeros(n_samples(len(c)), np.float32)
            ns += len(wav)
    torch.cuda.synchronize()
    w2 = time.perf_counter()
    sync_all()
    w3 = time.perf_counter()
    for dd in extra_dec:
        dd.close()
    nf = sum(len(c) for c in codes)
    wave = max(1, min(16, int(os.environ.get("Q3T_BATCH_WAVE", "16"))))     # sequences per lm_batch_kernel launch (engine.cu)
    groups = (len(pairs) + wave - 1) // wave
    stats = torch.tensor([w3 - w0, w1 - w0, w2 - w1, tm.prefill_ms * 1e-3, tm.frames_ms * 1e-3], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(ns), float(nf)], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    if rank != 0:
        return None
    step_s, gen_s, voc_s, prefill_s, frames_s = [float(x) for x in stats]
    ns_all, nf_all = float(tot[0]), float(tot[1])
    per_gpu = (n_utt + world - 1) // world
    # bytes one batched frame launch streams: the frame's weights once for all sequences + every sequence's KV history
    per_launch = min(wave, max(len(pairs), 1))
    frame_bytes = CODEPRED_FRAME_BYTES[args.wtype] + TALKER_STEP_BYTES[args.wtype] + per_launch * KV_BYTES_PER_POS * (10 + args.frames / 2.0 + 1)
    ms_frame = frames_s * 1e3 / max(groups * args.frames, 1)
    peak, _ = measured_peaks()
    return {"utterances": n_utt, "frames_each": args.frames, "n_gpus": world, "utterances_per_gpu": per_gpu, "sequences_per_launch": per_launch,
            "value": ns_all / 24000.0 / step_s, "unit": UNIT, "scaling": "strong", "ms_per_step": step_s * 1e3,
            "phases_s_max_over_ranks": {"prompt_prefill": prefill_s, "frames": frames_s, "generate_wall": gen_s, "vocoder": voc_s},
            "ms_per_batched_frame": ms_frame, "codes_only_audio_s_per_s": nf_all / FRAME_RATE / gen_s,
            "roofline": {"kernel": f"lm_batch_kernel<8> (one launch = one frame of {per_launch} sequences: two interleaved lock-step groups of <= 8)", "bound": "hbm",
                         "algorithmic_bytes_per_launch": int(frame_bytes), "achieved": frame_bytes / (ms_frame * 1e-3) / 1e9,
                         "peak": peak, "unit": "GB/s", "frac": frame_bytes / (ms_frame * 1e-3) / 1e9 / peak,
                         "note": "weights are read once per launch for all its sequences, so HBM is not the binding resource here: the launch is "
                                 "bound by the FMA pipe (16 x the batch-1 FMAs) and by the SM's ingest of the activation packets (DESIGN.md 4.1b)"},
            "limiter": (f"{per_gpu} utterances per GPU = {groups} launch wave(s) of <= {wave} sequences: a batched frame costs ~3.4 ms for 8 and "
                        f"~5.8 ms for 16 sequences, prompts/prefills ({prefill_s:.2f} s) and vocoder decodes ({voc_s:.2f} s) run one utterance at a "
                        f"time and do not shrink below one wave's share"),
            "vocoder_instances": args.decoders if have_voc else 0,
            "workload": "config 4: 64 utterances (8-64 byte texts, seed 4), greedy, host buffers in/out, utterance i -> rank i mod N"}


def run_sharded_workload(args, rank, world, local_rank, t, d, have_voc, dist, sync_all, m):
    """Config 4: N independent utterances per step, request-parallel over the ranks (no collective on
    the data path; torch.distributed only gathers the per-utterance summaries and the timing)."""
    import importlib
    import torch
    disp = importlib.import_module("q3t_b200.dispatch")
    utts = disp.make_workload(args.utterances, seed=4, max_len=args.frames)

    def synth(tokens, spk, max_len):
        codes = t.generate(tokens, spk, max_len, repetition_penalty=1.05, temperature=0.0, top_k=0)
        if codes is None:
            raise SystemExit("generate: " + t.get_error())
        wav = d.decode(codes) if have_voc else np.zeros(n_samples(len(codes)), np.float32)
        return codes, wav

    def synth_batch(mine):
        lens = {u.max_len for u in mine}
        if len(lens) != 1:
            raise SystemExit("config 4 uses one max_len for all utterances")
        codes = t.generate_batch([(u.tokens, u.spk) for u in mine], mine[0].max_len, repetition_penalty=1.05)
        if codes is None:
            raise SystemExit("generate_batch: " + t.get_error())
        return [(c, d.decode(c) if have_voc else np.zeros(n_samples(len(c)), np.float32)) for c in codes]

    batched = not args.no_batch
    run = (lambda: disp.run_sharded(utts, None, rank, world, dist, synth_batch=synth_batch)) if batched else \
          (lambda: disp.run_sharded(utts, synth, rank, world, dist))
    for _ in range(max(1, min(args.warmup, 1))):
        run()
    sampler = ClockSampler(local_rank)
    launches0 = t.timing().kernel_launches
    sync_all()
    sampler.start()
    w0 = time.perf_counter()
    for _ in range(args.steps):
        local, merged = run()
    sync_all()
    step_s = (time.perf_counter() - w0) / args.steps
    clocks = sampler.stop()
    times = torch.tensor([step_s], dtype=torch.float64, device="cuda")
    if dist:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    step_s = float(times[0])
    if rank == 0:
        audio_s = sum(v[1] for v in merged.values()) / 24000.0
        h2d = sum(u.tokens.nbytes + u.spk.nbytes for u in utts)
        d2h = sum(v[0] * 16 * 4 + v[1] * 4 for v in merged.values())
        emit_json(({
            "metric": METRIC, "value": audio_s / step_s, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": 1, "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f16", "data": "synthetic (seeded random-init GGUF of the real 0.6B shapes)",
            "config": {"workload": f"config 4: {args.utterances} utterances (8-64 byte texts, seed 4), greedy, {args.frames} frames each + vocoder, "
                                   f"utterance i -> rank i mod {world}, " + ("16 sequences per kernel launch on each GPU (lm_batch_kernel, two interleaved lock-step groups of 8)" if batched else "one utterance at a time per GPU (batch-1 engine)"),
                       "utterances": args.utterances, "frames": args.frames, "parallelism": f"request-parallel x{world} (no collective)",
                       "l2": "weights streamed per frame (3.3 GB) exceed the 126 MB L2; no explicit flush"},
            "e2e": {"value": audio_s / step_s, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
            "gpu_launches": int(t.timing().kernel_launches - launches0), "clocks": clocks,
        }))
    if dist:
        dist.barrier()
        dist.destroy_process_group()


_JSON_FD = None


def emit_json(obj):
    line = (json.dumps(obj) + "\n").encode()
    sys.stdout.flush()
    os.write(_JSON_FD if _JSON_FD is not None else 1, line)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=320)
    ap.add_argument("--ref-frames", type=int, default=0, help="frames per step of the CPU arms; 0 = --frames (the same workload)")
    ap.add_argument("--wtype", default="f16", choices=["f16", "q8_0"])
    ap.add_argument("--temperature", type=float, default=0.0, help="0 = greedy (configs 1/2/4/5); config 3 uses 0.9")
    ap.add_argument("--top-k", type=int, default=0, help="config 3 uses 50")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-vocoder", action="store_true", help="debug only: codes without the waveform (not a valid bench value)")
    ap.add_argument("--no-batch", action="store_true", help="config 4 through the batch-1 engine, one utterance at a time")
    ap.add_argument("--decoders", type=int, default=0, help="config 4: vocoder instances (host threads, one stream each) that decode the job's "
                                                            "utterances; 0 = min(4, host cores / (2 x ranks)), so 8 ranks do not oversubscribe the host")
    ap.add_argument("--utterances", type=int, default=0,
                    help="config 4: this many independent utterances (8-64 byte texts) per step, dealt i mod G to the ranks; "
                         "0 = the default config-2 workload (one utterance per rank per step)")
    args = ap.parse_args()
    # stdout carries exactly ONE line, the JSON: libraries that write to file descriptor 1 (the NCCL version banner at
    # communicator creation, compiler chatter) are sent to stderr, and the JSON goes to the saved descriptor
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
