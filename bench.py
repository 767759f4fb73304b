#!/usr/bin/env python
"""bench.py — audio-seconds per second of the Fun-ASR-Nano audio hot path (fbank+LFR+CMVN -> SAN-M encoder -> adaptor).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload cfg5|cfg2|cfg3] [--no-extras]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One "step" = one pass of the whole path over one batch of synthetic utterances (seeded, random-init weights of the named
architecture).  Workloads (BASELINE.json configs):
  cfg5 (default)  1 h of audio per GPU = 360 x 10 s utterances, one packed batch         (weak scaling: per-GPU work fixed)
  cfg2            16 x 10 s                                                              (CUDA-graph replay path)
  cfg3            64 utterances of 1-30 s (seed 0), packed offsets
  cfg1            one 10 s waveform through the frontend only (reported inside "workloads")
Prints ONE JSON line (rank 0).  `value` = device-resident throughput (inputs already in HBM), `e2e` = the same through
AudioPipeline with pinned HOST buffers (H2D of the PCM and D2H of the bf16 prompt-embed rows inside the timed region).
Besides the contract keys the line carries, at N = 1: "workloads" (cfg1 / cfg2 / cfg3 measured the same way), "e2e_dropin"
(the three reference calls extract_fbank -> audio_encoder -> audio_adaptor on CPU waveforms, infer.py:446-466), "gpu_eager"
(what a reference user gets on this GPU today: the plain PyTorch modules on cuda, fp32 and bf16) and "cpu_baseline"; at every
N: "strong" (ONE hour partitioned by shard_lpt over the N ranks, uniform and 1-30 s variable-length).
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
for p in (ROOT, os.path.join(ROOT, "fun-asr-vllm_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "audio-sec/sec (RTFx) fbank+LFR+CMVN+SAN-M encoder+adaptor"
UNIT = "audio-s/s"
ENC_CONF = dict(output_size=512, attention_heads=4, linear_units=2048, num_blocks=50, tp_blocks=20, kernel_size=11)
AD_CONF = dict(downsample_rate=1, encoder_dim=512, llm_dim=1024, ffn_dim=2048, n_layer=2, attention_heads=8)
GEMM_FLOP_PER_TOKEN = 465_715_200  # SURVEY.md §8d / BASELINE.md §4
ATTN_FLOP_PER_TOKEN_PER_T = 151_552
# Frontend: warp instructions the kernel executes per frame pair of the utterances (ncu "Instructions Executed" of a cfg5 launch,
# profiles/r02/ncu_source_frontend_r2bi.csv.gz: 255.6 M in all, 105.7 M of them fp64 (DADD / DFMA / DMUL), for 179 640 frame pairs;
# the tile overlap of one frame in 31 and the per-tile code are included).  A sub-partition issues one warp instruction per clock
# and its fp64 pipe takes one every two clocks.
FE_INSTR_PER_FRAME_PAIR = 1423
FE_FP64_INSTR_PER_FRAME_PAIR = 588
SM_SUBPARTITIONS = 148 * 4


def workload_lengths(name: str) -> np.ndarray:
    if name == "cfg5":
        return np.full(360, 160000, np.int64)
    if name == "cfg2":
        return np.full(16, 160000, np.int64)
    if name == "cfg3":
        return (np.random.default_rng(0).uniform(1.0, 30.0, 64) * 16000).astype(np.int64)
    if name == "cfg1":
        return np.full(1, 160000, np.int64)
    if name == "varlen_hour":  # 1-30 s utterances (seed 0) until one hour of audio is reached
        u = (np.random.default_rng(0).uniform(1.0, 30.0, 400) * 16000).astype(np.int64)
        return u[: int(np.searchsorted(np.cumsum(u), 3600 * 16000)) + 1]
    raise ValueError(name)


def synth_wave(n: int, seed: int):
    import torch

    g = torch.Generator().manual_seed(seed)
    return 0.5 * (2 * torch.rand(int(n), generator=g) - 1)  # SURVEY.md §8d set 1


def lfr_len(n: np.ndarray) -> np.ndarray:
    m = 1 + (n - 400) // 160
    return -(-m // 6)


def flops_of(T: np.ndarray) -> tuple[float, float]:
    T = T.astype(np.float64)
    return float(GEMM_FLOP_PER_TOKEN * T.sum()), float(ATTN_FLOP_PER_TOKEN_PER_T * (T * T).sum())


def random_init_(module, seed: int):
    """Random-init weights of the named architecture (no checkpoint on this box): U(+-sqrt(3/fan_in)) matrices,
    LN gamma ~ 1, small biases — the same recipe as oracle.encoder.seeded_init_ but product-side (no oracle import)."""
    import torch

    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in sorted(module.named_parameters()):
            u = torch.rand(p.shape, generator=g) * 2 - 1
            if "norm" in name:
                p.copy_(1.0 + 0.1 * u if name.endswith("weight") else 0.1 * u)
            elif name.endswith("bias"):
                p.copy_(0.1 * u)
            else:
                fan_in = p[0].numel()
                p.copy_(u / fan_in ** 0.5 * (3.0 ** 0.5 if "fsmn" not in name else 1.0))
    return module.eval()


def make_cmvn() -> np.ndarray:
    rng = np.random.default_rng(99)
    return np.stack([-(8 + 4 * rng.random(560)), 0.1 + 0.2 * rng.random(560)]).astype(np.float32)


# ----------------------------------------------------------------------------- clocks sampler
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0: float, t1: float) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for ts, r in self.rows if t0 - 0.05 <= ts <= t1 + 0.15] or [r for _, r in self.rows]
        sm, mx, reasons = [], None, set()
        for r in rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------- CPU reference path
# The reference's own CPU arithmetic for this path, timed on the host cores.  Frontend leg: the on-box copy of
# funasr's WavFrontend / extract_fbank (vLLM ships it verbatim) on the REAL torchaudio kaldi.fbank when importable
# ("reference"), else the numpy restatement in oracle/ ("port").  Encoder + adaptor leg: the fp32 PyTorch restatement in
# oracle/ ("port": funasr itself is not installable here, SURVEY.md §0, and vLLM's classes cannot be built outside an engine).
# This is the ONLY use of oracle/ outside tests/ and smoke() (plus the informational gpu_eager leg below).
class CpuReference:
    def __init__(self, cmvn: np.ndarray):
        import torch

        from oracle import encoder as oe

        self.torch = torch
        self.cmvn = cmvn
        self.enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **oe.NANO_ENCODER_CONF), seed=4321)
        self.ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)
        self.all_cores = len(os.sched_getaffinity(0))
        self.frontend, self.extract_fbank, self.frontend_kind = None, None, "port"
        try:
            from vllm.transformers_utils.processors.funasr import FunASRFeatureExtractor, WavFrontend

            extract_fbank = FunASRFeatureExtractor().extract_fbank  # funasr's extract_fbank, a method in vLLM's copy (VLLM:292-317)
            fe = WavFrontend(cmvn_file=None, fs=16000, window="povey", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6,
                             dither=0.0)
            fe.cmvn = torch.from_numpy(cmvn)  # synthetic am.mvn values (the parser is not what is timed)
            self.frontend, self.extract_fbank, self.frontend_kind = fe, extract_fbank, "reference"
        except Exception:
            pass

    def step(self, wavs) -> tuple[float, float]:
        """One pass of the reference's batched loop (infer.py:446-458) on CPU; returns (frontend s, encoder+adaptor s)."""
        torch = self.torch
        t0 = time.perf_counter()
        if self.frontend is not None:
            feats, lens = self.extract_fbank(wavs, frontend=self.frontend, is_final=True)  # infer.py:446
            lens = lens.long()
        else:
            from oracle import frontend as of

            f, l = of.wav_frontend([w.numpy() for w in wavs], "povey", self.cmvn)
            feats, lens = torch.from_numpy(f), torch.from_numpy(l).long()
        t1 = time.perf_counter()
        with torch.inference_mode():
            y, _ = self.enc.forward_padded(feats, lens)  # infer.py:455 on the padded batch
            self.ad._run(y)  # infer.py:458
        return t1 - t0, time.perf_counter() - t1

    def timed(self, wavs, threads: int, repeats: int = 5, warmup: int = 2) -> dict:
        self.torch.set_num_threads(threads)
        for _ in range(warmup):
            self.step(wavs[:1])
        best = None
        for _ in range(repeats):
            f, e = self.step(wavs)
            if best is None or f + e < best[0] + best[1]:
                best = (f, e)
        secs = float(sum(int(w.numel()) for w in wavs)) / 16000.0
        return {"threads": threads, "utterances": len(wavs), "audio_s": secs, "frontend_s": best[0], "encoder_adaptor_s": best[1],
                "frontend_audio_s_per_s": secs / best[0], "encoder_adaptor_audio_s_per_s": secs / best[1],
                "audio_s_per_s": secs / (best[0] + best[1])}


def cpu_baseline(lengths: np.ndarray, cmvn: np.ndarray) -> dict:
    """warm-up 2, best of 5, 1 thread and all cores, on a bounded sample of the workload."""
    ref = CpuReference(cmvn)
    idx = np.linspace(0, len(lengths) - 1, min(16, len(lengths))).astype(int)
    wavs = [synth_wave(lengths[i], 1000 + int(i)) for i in idx]
    one = ref.timed(wavs[:1], 1, repeats=3, warmup=1)
    full = ref.timed(wavs, ref.all_cores)
    fe_leg = "real WavFrontend + torchaudio.compliance.kaldi.fbank" if ref.frontend_kind == "reference" else "numpy port"
    return {"value": full["audio_s_per_s"], "unit": UNIT, "cores": ref.all_cores, "kind": "port", "frontend_kind": ref.frontend_kind,
            "sample": f"{len(wavs)} utterances of the workload ({full['audio_s']:.0f} s audio) as one padded fp32 batch (infer.py:446-458), "
                      f"warm-up 2, best of 5; frontend leg = {fe_leg}, encoder/adaptor leg = fp32 PyTorch port",
            "all_cores": full, "one_thread": one}


def run_reference(args, rank: int) -> None:
    if rank != 0:
        return
    lengths = workload_lengths(args.workload)
    ref = CpuReference(make_cmvn())
    n = min(16, len(lengths))  # the reference's own batch size (infer.py:415 batch_size=16)
    idx = np.linspace(0, len(lengths) - 1, n).astype(int)
    wavs = [synth_wave(lengths[i], 1000 + int(i)) for i in idx]
    secs = float(sum(lengths[i] for i in idx)) / 16000.0
    ref.torch.set_num_threads(ref.all_cores)
    for _ in range(args.warmup):
        ref.step(wavs[:2])  # bounded warm-up: 2 utterances
    ts = [sum(ref.step(wavs)) for _ in range(args.steps)]
    t = float(np.min(ts))  # best of the timed steps
    val = secs / t
    sample = (f"each step = {n} utterances of {args.workload} ({secs:.0f} s audio) as one padded fp32 batch (infer.py:446-458); best of "
              f"{args.steps} steps; frontend leg: {ref.frontend_kind}, encoder/adaptor leg: port")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": {"workload": args.workload, "sample": sample, "device": "host CPU",
                                                         "ms_per_step_mean": 1e3 * float(np.mean(ts))},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": ref.all_cores, "kind": "port", "frontend_kind": ref.frontend_kind,
                         "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ----------------------------------------------------------------------------- B200 arm
class Inputs:
    """Seeded synthetic utterances of one workload: CPU waveforms, the packed pinned host PCM and its device copy."""

    def __init__(self, lengths: np.ndarray, seed0: int, dev, llm_dim: int):
        import torch

        self.lengths = lengths
        self.T = lfr_len(lengths)
        self.rows = int(self.T.sum())
        self.audio_s = float(lengths.sum()) / 16000.0
        base = [synth_wave(int(lengths.max()), seed0 + i) for i in range(min(8, len(lengths)))]  # re-used cyclically
        self.wavs = [base[i % len(base)][: int(n)] for i, n in enumerate(lengths)]
        self.starts = np.zeros(len(lengths), np.int64)
        self.starts[1:] = np.cumsum((lengths[:-1] + 3) // 4 * 4)
        total = int(self.starts[-1] + (lengths[-1] + 3) // 4 * 4)
        self.host_pcm = torch.zeros(total, dtype=torch.float32).pin_memory()
        for w, s0, n in zip(self.wavs, self.starts, lengths):
            self.host_pcm[s0:s0 + n] = w
        self.dev_pcm = self.host_pcm.to(dev)
        self.out_dev = torch.empty((self.rows, llm_dim), dtype=torch.bfloat16, device=dev)
        self.out_host = torch.empty((self.rows, llm_dim), dtype=torch.bfloat16).pin_memory()


def run_b200(args, rank: int, local_rank: int, world: int) -> None:
    import torch
    import torch.distributed as dist

    import funasr_b200 as fb
    from funasr_b200 import _lib
    from funasr_b200.sharding import shard_lpt

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    numa_cores = None
    if world > 1:
        from funasr_b200.sharding import bind_to_gpu_numa

        numa_cores = bind_to_gpu_numa(local_rank)  # pinned staging buffers on the GPU's own NUMA node
        # NCCL prints its version banner on fd 1 when the communicator comes up; stdout carries ONE JSON line, so the banner
        # goes to stderr
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(*vals):
        if world == 1:
            return list(vals)
        t = torch.tensor(vals, device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()

    # ---- model (random init of the named architecture) and inputs
    enc = random_init_(fb.SenseVoiceEncoderSmall(560, **ENC_CONF), 4321).to(dev)
    ad = random_init_(fb.Transformer(**AD_CONF), 4322).to(dev)
    cmvn = make_cmvn()
    fe = fb.WavFrontend(fs=16000, window="povey", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                        cmvn=torch.from_numpy(cmvn), device=dev)
    main_in = Inputs(workload_lengths(args.workload), 1000 + 97 * rank, dev, AD_CONF["llm_dim"])
    tower = fb.AudioTower(enc, ad, max_tokens=max(main_in.rows, 61440))
    pipe = fb.AudioPipeline(fe, tower, dev)
    graph_rows = args.graph_rows if args.graph_rows is not None else 8192
    n_streams = args.streams if args.streams is not None else 2
    tower.plan().set_graph_rows(graph_rows)
    tower.plan().set_streams(n_streams)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def timed(fn, k):
        """K steps, each bracketed by CUDA events on the launching stream; L2 flushed (untimed) before every step."""
        evs = []
        for _ in range(k):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            evs.append((a, b))
        torch.cuda.synchronize()
        return sum(a.elapsed_time(b) for a, b in evs) / k

    def steps_of(inp: Inputs):
        def step_device():
            pipe.embed_device(inp.dev_pcm, inp.starts, inp.lengths, out=inp.out_dev)

        def step_e2e():
            pipe.run_pinned(inp.host_pcm, inp.starts, inp.lengths, inp.out_host, overlap=False if args.no_overlap else None,
                            pieces=args.e2e_pieces)

        return step_device, step_e2e

    def profile(inp: Inputs):
        """launches per step + per-kernel-class device time (profiled steps, NOT the timed ones; launch by launch on ONE stream:
        event pairs around launches that overlap on two streams would double count, and a replayed graph has no brackets)."""
        step_device, _ = steps_of(inp)
        tower.plan().set_streams(1)
        _lib.profile_enable(True)
        for _ in range(2):
            flush.zero_()
            step_device()
        prof = _lib.profile_collect()
        _lib.profile_enable(False)
        tower.plan().set_streams(n_streams)
        step_device()
        return {k: (ms / 2, n // 2) for k, (ms, n) in prof.items()}

    def measure(inp: Inputs, steps: int, warmup: int, sync=torch.cuda.synchronize):
        step_device, step_e2e = steps_of(inp)
        for _ in range(warmup):
            step_device()
        sync()
        ms_dev = timed(step_device, steps)
        sync()
        for _ in range(2):
            step_e2e()
        sync()
        ms_e2e = timed(step_e2e, steps)
        sync()
        return ms_dev, ms_e2e

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_tf, peak_src = (peaks.get("bf16_tflops_sustained"), "measured sustained") if peaks else (1400.0, "fallback sustained")
    hbm = peaks.get("hbm_gbs", 6650.0)

    def gemm_roofline(inp: Inputs, prof, ms_step):
        gemm_flop, _ = flops_of(inp.T)
        gemm_ms, gemm_n = prof["gemm"]
        achieved = gemm_flop / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
        share = gemm_ms * ms_step / max(sum(ms for ms, _ in prof.values()), 1e-9)
        return achieved, gemm_ms, gemm_n, (gemm_flop / (share * 1e-3) / 1e12 if share > 0 else 0.0)

    # ---- main workload: warm-up, profile, then the timed region (barrier + synchronize on both sides, max over ranks)
    warm = max(args.warmup, 3)
    step_device, step_e2e = steps_of(main_in)
    for _ in range(warm):
        step_device()
    torch.cuda.synchronize()
    prof = profile(main_in)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    barrier()
    t0 = time.time()
    ms_dev = timed(step_device, args.steps)
    barrier()
    t1 = time.time()
    for _ in range(2):
        step_e2e()
    barrier()
    ms_e2e = timed(step_e2e, args.steps)
    barrier()
    clocks = sampler.stop(t0, t1) if sampler else None
    # the same end-to-end step from int16 host PCM (the sample type of a 16 kHz wav file): half the H2D bytes, x / 32768 on the GPU
    host_pcm16 = (main_in.host_pcm * 32767.0).round().to(torch.int16).pin_memory()

    def step_e2e16():
        pipe.run_pinned(host_pcm16, main_in.starts, main_in.lengths, main_in.out_host, overlap=False if args.no_overlap else None,
                        pieces=args.e2e_pieces)

    for _ in range(2):
        step_e2e16()
    barrier()
    ms_e2e16 = timed(step_e2e16, args.steps)
    barrier()
    # the same end-to-end step in a serving loop: batches submitted back to back to two lanes (own streams, workspaces and buffers),
    # so that the copies of one batch run under the compute of another; every step still makes its own H2D and D2H
    sp = fb.StreamingPipeline(fe, pipe.tower, dev)
    outs = [main_in.out_host, torch.empty_like(main_in.out_host).pin_memory()]
    n_sp = [0]

    def step_stream():
        sp.submit(main_in.host_pcm, main_in.starts, main_in.lengths, outs[n_sp[0] % 2], overlap=False if args.no_overlap else None)
        n_sp[0] += 1

    def timed_stream(k):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(k):
            step_stream()
        for st in sp.streams:
            torch.cuda.current_stream().wait_stream(st)
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / k

    timed_stream(2)
    barrier()
    ms_stream = timed_stream(max(args.steps, 4))
    barrier()
    ms_dev, ms_e2e, ms_e2e16, ms_stream = max_over_ranks(ms_dev, ms_e2e, ms_e2e16, ms_stream)

    # ---- strong scaling: ONE hour of audio partitioned by shard_lpt over the ranks (every rank computes the same partition)
    strong = {}
    if not args.no_extras:
        for name in ("cfg5", "varlen_hour"):
            lengths = workload_lengths(name)
            T_all = lfr_len(lengths)
            parts = shard_lpt(T_all.tolist(), world)
            mine = Inputs(lengths[np.asarray(parts[rank], np.int64)], 5000, dev, AD_CONF["llm_dim"])
            d, e = measure(mine, max(3, args.steps // 2), 2, sync=barrier)
            d, e = max_over_ranks(d, e)
            hour_s = float(lengths.sum()) / 16000.0
            if world == 1:
                d1, e1 = d, e
            else:  # the whole hour on ONE GPU, measured on rank 0 in this very run (the others wait at the barrier)
                d1 = e1 = 0.0
                if rank == 0:
                    d1, e1 = measure(Inputs(lengths, 5000, dev, AD_CONF["llm_dim"]), max(3, args.steps // 2), 2)
                barrier()
                d1, e1 = max_over_ranks(d1, e1)
            strong["uniform_hour" if name == "cfg5" else name] = {
                "utterances": int(len(lengths)), "audio_s": hour_s,
                "rows_per_rank_max": int(max(T_all[np.asarray(p, np.int64)].sum() for p in parts)),
                "ms": d, "audio_s_per_s": hour_s / (d * 1e-3), "e2e_ms": e, "e2e_audio_s_per_s": hour_s / (e * 1e-3),
                "one_gpu_ms": d1, "one_gpu_e2e_ms": e1, "efficiency": d1 / (world * d), "e2e_efficiency": e1 / (world * e)}

    if rank == 0:
        achieved, gemm_ms, gemm_n, achieved_share = gemm_roofline(main_in, prof, ms_dev)
        # DRAM traffic per launch from the committed `ncu --set full` capture (profiles/ncu_traffic.json; bytes, cfg5-sized
        # launches only: the capture ran each kernel alone on 60 120 rows)
        traffic_gemm = traffic_fe = None
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
            if args.workload == "cfg5":
                traffic_gemm = tr["gemm_mean_layer_launch_MB"] * 1e6
                traffic_fe = (tr["kernels"]["frontend_1h"]["dram_read_MB"] + tr["kernels"]["frontend_1h"]["dram_write_MB"]) * 1e6
        except Exception:
            pass
        rows = main_in.rows
        line = {
            "metric": METRIC, "value": world * main_in.audio_s / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": args.workload, "utterances_per_gpu": int(len(main_in.lengths)), "audio_s_per_gpu": main_in.audio_s,
                       "tokens_per_gpu": rows, "model": "SenseVoice SAN-M 50+20 blocks d512 ff2048 + adaptor 2x(d1024,8h), random init",
                       "frontend": "povey, 80 mel, LFR 7/6, CMVN", "l2": "256 MiB flush before every timed step; weights 467 MB > L2",
                       "sharding": "by utterance, no collective",
                       "host_binding": (f"rank bound to {len(numa_cores)} GPU-local cores" if numa_cores else "none"),
                       "graph_rows": graph_rows, "streams": n_streams,
                       "e2e_schedule": "single pass" if (args.no_overlap or rows < 8192 or len(main_in.lengths) < 8) else
                       "two half-batches on two streams: H2D of half 2 and D2H of half 1 under compute"},
            "e2e": {"value": world * main_in.audio_s / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(main_in.host_pcm.numel() * 4), "d2h_bytes_per_step": int(main_in.out_host.numel() * 2)},
            "e2e_pcm16": {"value": world * main_in.audio_s / (ms_e2e16 * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e16,
                          "h2d_bytes_per_step": int(host_pcm16.numel() * 2), "d2h_bytes_per_step": int(main_in.out_host.numel() * 2),
                          "note": "informational: int16 host PCM, converted on the GPU (fa_pcm_to_mono_f32); the headline e2e is the f32 one"},
            "e2e_streaming": {"value": world * main_in.audio_s / (ms_stream * 1e-3), "unit": UNIT, "ms_per_step": ms_stream,
                              "h2d_bytes_per_step": int(main_in.host_pcm.numel() * 4), "d2h_bytes_per_step": int(main_in.out_host.numel() * 2),
                              "note": "informational: StreamingPipeline, batches back to back on two lanes, no L2 flush between steps; "
                                      "the headline e2e is one batch at a time"},
            "gpu_launches": int(sum(n for _, n in prof.values())),
            "roofline": {"bound": "tensor", "kernel": "gemm_bf16_kernel (tcgen05, all 70x4+10 GEMMs of a step)", "achieved": achieved,
                         "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved / peak_tf if peak_tf else None, "traffic": traffic_gemm,
                         "peak_source": peak_src, "launches": gemm_n, "ms_per_step": gemm_ms,
                         # the profiled step brackets every launch with an event pair, which serialises the launches
                         # (no programmatic-dependent-launch overlap); the same class time scaled to the timed step:
                         "achieved_share_of_timed_step": achieved_share},
            "kernel_ms_per_step": {k: round(ms, 4) for k, (ms, n) in prof.items()},
            "kernel_launches_per_step": {k: n for k, (ms, n) in prof.items()},
            "clocks": clocks,
        }
        fe_ms = prof["frontend"][0]
        if fe_ms > 0:
            fe_bytes = float(4 * main_in.lengths.sum() + 4 * 560 * main_in.T.sum())
            frames = float((1 + (main_in.lengths - 400) // 160).sum())
            fe_clk = 1e6 * float((clocks or {}).get("sm_max_mhz") or 1965.0)
            line["roofline_frontend"] = {"bound": "hbm", "kernel": "fbank_kernel<0>", "achieved": fe_bytes / (fe_ms * 1e-3) / 1e9,
                                         "peak": hbm, "unit": "GB/s", "frac": fe_bytes / (fe_ms * 1e-3) / 1e9 / hbm, "traffic": traffic_fe,
                                         "ms": fe_ms,
                                         # the FFT runs in fp64 (parity): share of the fp64 pipe's and of the
                                         # issue slots' cycles the kernel's instruction counts account for, at the max SM clock
                                         "frac_fp64": frames / 2 * FE_FP64_INSTR_PER_FRAME_PAIR * 2 / (SM_SUBPARTITIONS * fe_clk * fe_ms * 1e-3),
                                         "frac_issue": frames / 2 * FE_INSTR_PER_FRAME_PAIR / (SM_SUBPARTITIONS * fe_clk * fe_ms * 1e-3),
                                         "sm_clock_hz_assumed": fe_clk}
        if strong:
            line["strong"] = strong
        if world == 1 and not args.no_extras:
            ctx = dict(args=args, fb=fb, pipe=pipe, fe=fe, dev=dev, measure=measure, profile=profile, gemm_roofline=gemm_roofline,
                       timed=timed, peak_tf=peak_tf, hbm=hbm, graph_rows=graph_rows)
            line["workloads"] = extra_workloads(**ctx)
            line["e2e_dropin"] = e2e_dropin(fb, fe, enc, ad, dev)
            try:
                line["gpu_eager"] = gpu_eager(dev)
            except Exception as ex:  # informational leg: never fails the bench
                line["gpu_eager"] = {"unavailable": repr(ex)[:200]}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(main_in.lengths, cmvn)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def extra_workloads(args, fb, pipe, fe, dev, measure, profile, gemm_roofline, timed, peak_tf, hbm, graph_rows) -> dict:
    """cfg1 / cfg2 / cfg3 of BASELINE.json measured exactly like the main workload (device-resident + e2e, L2 flushed)."""
    import torch

    out = {}
    for name in ("cfg2", "cfg3"):
        if name == args.workload:
            continue
        inp = Inputs(workload_lengths(name), 2000, dev, AD_CONF["llm_dim"])
        d, e = measure(inp, 20, 5)
        prof = profile(inp)
        achieved, gemm_ms, gemm_n, share = gemm_roofline(inp, prof, d)
        gemm_flop, attn_flop = flops_of(inp.T)
        out[name] = {"utterances": int(len(inp.lengths)), "rows": inp.rows, "audio_s": inp.audio_s, "ms_per_step": d,
                     "audio_s_per_s": inp.audio_s / (d * 1e-3), "e2e_ms": e, "e2e_audio_s_per_s": inp.audio_s / (e * 1e-3),
                     "schedule": "CUDA-graph replay" if inp.rows <= graph_rows else "dual stream",
                     "gemm_tflops_launch_by_launch": achieved, "gemm_frac": achieved / peak_tf,
                     "gemm_tflops_share_of_timed_step": share, "gemm_frac_share_of_timed_step": share / peak_tf,
                     "whole_step_tflops": (gemm_flop + attn_flop) / (d * 1e-3) / 1e12,
                     "kernel_ms_launch_by_launch": {k: round(ms, 4) for k, (ms, n) in prof.items()}}
        # a serving loop over batches of this size (StreamingPipeline: batches back to back on 2 / 4 lanes, every batch with its own
        # H2D and D2H, no L2 flush between them): what concurrency gives back where one batch alone cannot fill the GPU
        for lanes in (2, 4):
            sp = fb.StreamingPipeline(fe, pipe.tower, dev, lanes=lanes)
            outs = [torch.empty_like(inp.out_host).pin_memory() for _ in range(lanes)]

            def run(k):
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for i in range(k):
                    sp.submit(inp.host_pcm, inp.starts, inp.lengths, outs[i % lanes])
                for st in sp.streams:
                    torch.cuda.current_stream().wait_stream(st)
                b.record()
                torch.cuda.synchronize()
                return a.elapsed_time(b) / k

            run(2 * lanes)
            ms = run(40)
            out[name][f"e2e_streaming_{lanes}_lanes_ms"] = ms
            out[name][f"e2e_streaming_{lanes}_lanes_audio_s_per_s"] = inp.audio_s / (ms * 1e-3)
            del sp, outs
    # cfg1: fbank + LFR + CMVN on one 10 s waveform, batch 1 (launch/latency bound: 1 MB of traffic)
    inp = Inputs(workload_lengths("cfg1"), 3000, dev, AD_CONF["llm_dim"])
    for _ in range(5):
        pipe.features(inp.dev_pcm, inp.starts, inp.lengths)
    d = timed(lambda: pipe.features(inp.dev_pcm, inp.starts, inp.lengths), 50)

    def fe_e2e():
        return fb.extract_fbank([inp.wavs[0]], frontend=fe, is_final=True)  # the reference call, CPU waveform in

    for _ in range(5):
        fe_e2e()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(50):
        fe_e2e()
    torch.cuda.synchronize()
    e = (time.perf_counter() - t0) / 50 * 1e3
    nbytes = float(4 * inp.lengths.sum() + 4 * 560 * inp.T.sum())
    out["cfg1"] = {"utterances": 1, "audio_s": inp.audio_s, "ms_per_step": d, "audio_s_per_s": inp.audio_s / (d * 1e-3),
                   "gbps": nbytes / (d * 1e-3) / 1e9, "hbm_frac": nbytes / (d * 1e-3) / 1e9 / hbm,
                   "e2e_ms_extract_fbank_wall": e, "e2e_audio_s_per_s": inp.audio_s / (e * 1e-3),
                   "note": "one launch over 1 MB: launch/latency bound, the HBM fraction is quoted on the cfg5-sized batch"}
    return out


def e2e_dropin(fb, fe, enc, ad, dev) -> dict:
    """The three reference calls on a list of CPU waveforms, as the reference's batched loop makes them (infer.py:446-466):
    extract_fbank -> audio_encoder -> audio_adaptor -> per-utterance slices, the padded fp32 result copied to the host.  The
    host pack loop, the padded fp32 intermediates, the repack kernels and the lens sync are all inside the timed region
    (wall clock, synchronised on both sides)."""
    import torch

    enc._shared_plan = ad._shared_plan = None  # the separate-call path: each module plans for itself
    out = {}
    for name in ("cfg2", "cfg5"):
        inp = Inputs(workload_lengths(name), 4000, dev, AD_CONF["llm_dim"])
        state = {"host": None}

        def step():
            speech, lens = fb.extract_fbank(inp.wavs, frontend=fe, is_final=True)        # infer.py:446
            speech, lens = speech.to(dev), lens.to(dev)                                 # infer.py:452-453
            y, ol = enc(speech, lens)                                                   # infer.py:455
            z, zl = ad(y, ol)                                                           # infer.py:458
            if state["host"] is None or state["host"].shape != z.shape:
                state["host"] = torch.empty(z.shape, dtype=z.dtype).pin_memory()
            state["host"].copy_(z, non_blocking=True)                                   # vLLM's renderer moves every tensor to the CPU
            return [state["host"][i, : int(n)] for i, n in enumerate(inp.T)]            # infer.py:463-466 slices

        for _ in range(3):
            step()
        torch.cuda.synchronize()
        ts = []
        for _ in range(5 if name == "cfg5" else 20):
            t0 = time.perf_counter()
            step()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        t = float(np.median(ts))
        out[name] = {"utterances": int(len(inp.lengths)), "ms_per_step": 1e3 * t, "audio_s_per_s": inp.audio_s / t,
                     "h2d_bytes_per_step": int(inp.host_pcm.numel() * 4), "d2h_bytes_per_step": int(state["host"].numel() * 4)}
    out["note"] = "wall clock, median; padded [B,Tmax,.] fp32 at every seam exactly as the reference passes it"
    return out


def gpu_eager(dev) -> dict:
    """Informational (SURVEY.md §2.3 'the Blackwell kernel to beat'): the reference's encoder + adaptor as plain PyTorch modules
    on this GPU — fp32 (the reference default, model.py:483-486) and bf16 (its opt-in) — i.e. cuBLAS/ATen eager on padded
    batches, device-resident features in, no frontend (the reference's frontend stays on the CPU).  Bench-only: uses the
    oracle's module definitions, which nothing under funasr_b200/ may import."""
    import torch

    from oracle import encoder as oe

    res = {}
    o_enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **oe.NANO_ENCODER_CONF), seed=4321)
    o_ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)

    def forward(e, a, x, lens, dtype):
        T = x.shape[1]
        mask = (torch.arange(T, device=dev)[None, :] < lens[:, None]).to(dtype)
        h = x * 512 ** 0.5 + oe.sinusoidal_pe(T, 560, torch.float32).to(dev)[None]  # PE in fp32 (bf16 PE is wrong)
        h = h.to(dtype)
        for l in list(e.encoders0) + list(e.encoders):
            h = l(h, mask)
        h = e.after_norm(h)
        for l in e.tp_encoders:
            h = l(h, mask)
        return a._run(e.tp_norm(h))

    for name in ("cfg2", "cfg5"):
        lengths = workload_lengths(name)
        T = lfr_len(lengths)
        x = torch.randn(len(T), int(T.max()), 560, device=dev)
        lens = torch.as_tensor(T, device=dev)
        audio_s = float(lengths.sum()) / 16000.0
        for dtype, tag in ((torch.float32, "fp32"), (torch.bfloat16, "bf16")):
            e, a = o_enc.to(dev, dtype), o_ad.to(dev, dtype)
            with torch.inference_mode():
                for _ in range(2):
                    forward(e, a, x, lens, dtype)
                torch.cuda.synchronize()
                ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(3)]
                for s0, s1 in ev:
                    s0.record()
                    forward(e, a, x, lens, dtype)
                    s1.record()
                torch.cuda.synchronize()
            ms = min(s0.elapsed_time(s1) for s0, s1 in ev)
            res[f"{name}_{tag}"] = {"ms_per_step": ms, "audio_s_per_s": audio_s / (ms * 1e-3)}
    res["note"] = ("encoder + adaptor only, padded batch resident on the GPU, torch eager (cuBLAS/ATen), best of 3; compare with "
                   "kernel_ms_per_step minus the frontend")
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg5", choices=["cfg5", "cfg2", "cfg3"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the workloads / e2e_dropin / gpu_eager / strong sections")
    ap.add_argument("--e2e-pieces", type=int, default=2, help="e2e leg: pieces of the overlapped schedule (two streams)")
    ap.add_argument("--no-overlap", action="store_true", help="e2e leg: one pass instead of the two-half overlapped schedule")
    ap.add_argument("--graph-rows", type=int, default=None, help="batches up to this many rows replay a CUDA graph (0 = off)")
    ap.add_argument("--streams", type=int, default=None, help="1 = single stream, 2 = dual-stream half-batch schedule")
    args = ap.parse_args()
    rank, local_rank, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if world != args.gpus:
        if args.gpus > 1 and world == 1:
            # convenience: re-launch under torchrun the way the driver does
            port = 29500 + os.getpid() % 2000
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}", "--master-addr",
                   "127.0.0.1", "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
            sys.exit(subprocess.call(cmd))
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    run_b200(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
