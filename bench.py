#!/usr/bin/env python
"""bench.py — audio-seconds per second of the Fun-ASR-Nano audio hot path (fbank+LFR+CMVN -> SAN-M encoder -> adaptor).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload cfg5|cfg2|cfg3]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P bench.py --gpus N ...

One "step" = one pass of the whole path over one batch of synthetic utterances (seeded, random-init weights of the named
architecture).  Workloads (BASELINE.json configs):
  cfg5 (default)  1 h of audio per GPU = 360 x 10 s utterances, one packed batch         (weak scaling: per-GPU work fixed)
  cfg2            16 x 10 s
  cfg3            64 utterances of 1-30 s (seed 0), packed offsets
Prints ONE JSON line (rank 0).  `value` = device-resident throughput (inputs already in HBM), `e2e` = the same through
AudioPipeline with pinned HOST buffers (H2D of the PCM and D2H of the bf16 prompt-embed rows inside the timed region).
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


def workload_lengths(name: str, rank: int = 0) -> np.ndarray:
    if name == "cfg5":
        return np.full(360, 160000, np.int64)
    if name == "cfg2":
        return np.full(16, 160000, np.int64)
    if name == "cfg3":
        return (np.random.default_rng(0).uniform(1.0, 30.0, 64) * 16000).astype(np.int64)
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


# ----------------------------------------------------------------------------- CPU reference path (oracle)
def cpu_reference_setup(n_threads: int | None = None):
    """The reference's own CPU arithmetic for this path, timed on the host cores: numpy restatement of
    torchaudio.kaldi.fbank + apply_lfr + apply_cmvn and the fp32 PyTorch restatement of SenseVoiceEncoderSmall + adaptor
    (oracle/, 'port': funasr itself is not installable here).  The ONLY use of oracle/ outside tests."""
    import torch

    from oracle import encoder as oe

    cores = n_threads or len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **oe.NANO_ENCODER_CONF), seed=4321)
    ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)
    return enc, ad, torch.get_num_threads()


def cpu_reference_step(enc, ad, wavs, cmvn) -> float:
    """One pass of the reference's batched loop (infer.py:446-458) on CPU; returns seconds."""
    import torch

    from oracle import frontend as of

    t0 = time.perf_counter()
    feats, lens = of.wav_frontend([w.numpy() for w in wavs], "povey", cmvn)  # python loop over utterances, like VLLM:137
    with torch.inference_mode():
        y, ol = enc.forward_padded(torch.from_numpy(feats), torch.from_numpy(lens).long())
        ad._run(y)
    return time.perf_counter() - t0


def cpu_baseline(lengths: np.ndarray, cmvn: np.ndarray, budget_s: float = 15.0) -> dict:
    enc, ad, cores = cpu_reference_setup()
    order = np.argsort(-lengths)[:64]
    probe = [synth_wave(lengths[order[0]], 777)]
    cpu_reference_step(enc, ad, probe, cmvn)  # warm-up (thread pools, allocator)
    t1 = cpu_reference_step(enc, ad, probe, cmvn)
    n = int(max(1, min(16, len(lengths), budget_s / max(t1, 1e-3))))
    idx = np.linspace(0, len(lengths) - 1, n).astype(int)
    wavs = [synth_wave(lengths[i], 1000 + int(i)) for i in idx]
    t = min(cpu_reference_step(enc, ad, wavs, cmvn) for _ in range(2))
    secs = float(sum(lengths[i] for i in idx)) / 16000.0
    return {"value": secs / t, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n} utterance(s) of the workload ({secs:.0f} s audio) as one padded fp32 batch, best of 2, {t:.2f} s"}


def run_reference(args, rank: int) -> None:
    if rank != 0:
        return
    lengths = workload_lengths(args.workload)
    cmvn = make_cmvn()
    enc, ad, cores = cpu_reference_setup()
    n = min(16, len(lengths))  # the reference's own batch size (infer.py:415 batch_size=16)
    idx = np.linspace(0, len(lengths) - 1, n).astype(int)
    wavs = [synth_wave(lengths[i], 1000 + int(i)) for i in idx]
    secs = float(sum(lengths[i] for i in idx)) / 16000.0
    for _ in range(args.warmup):
        cpu_reference_step(enc, ad, wavs[:2], cmvn)  # bounded warm-up: 2 utterances
    ts = [cpu_reference_step(enc, ad, wavs, cmvn) for _ in range(args.steps)]
    t = float(np.mean(ts))
    val = secs / t
    sample = f"each step = {n} utterances of {args.workload} ({secs:.0f} s audio) as one padded fp32 batch (infer.py:446-458)"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": {"workload": args.workload, "sample": sample, "device": "host CPU"},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def make_cmvn() -> np.ndarray:
    rng = np.random.default_rng(99)
    return np.stack([-(8 + 4 * rng.random(560)), 0.1 + 0.2 * rng.random(560)]).astype(np.float32)


# ----------------------------------------------------------------------------- B200 arm
def run_b200(args, rank: int, local_rank: int, world: int) -> None:
    import torch
    import torch.distributed as dist

    import funasr_b200 as fb
    from funasr_b200 import _lib

    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    numa_cores = None
    if world > 1:
        from funasr_b200.sharding import bind_to_gpu_numa

        numa_cores = bind_to_gpu_numa(local_rank)  # pinned staging buffers on the GPU's own NUMA node
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- model (random init of the named architecture) and inputs
    enc = random_init_(fb.SenseVoiceEncoderSmall(560, **ENC_CONF), 4321).to(dev)
    ad = random_init_(fb.Transformer(**AD_CONF), 4322).to(dev)
    cmvn = make_cmvn()
    fe = fb.WavFrontend(fs=16000, window="povey", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                        cmvn=torch.from_numpy(cmvn), device=dev)
    lengths = workload_lengths(args.workload, rank)
    T = lfr_len(lengths)
    audio_s = float(lengths.sum()) / 16000.0
    tower = fb.AudioTower(enc, ad, max_tokens=int(T.sum()))
    pipe = fb.AudioPipeline(fe, tower, dev)
    if args.chunk_tokens is not None:
        tower.plan().set_chunk_tokens(args.chunk_tokens)
    if args.streams is not None:
        tower.plan().set_streams(args.streams)
    # distinct seeded noise per utterance and rank (a few base waves re-used cyclically keep host set-up short)
    base = [synth_wave(int(lengths.max()), 1000 + 97 * rank + i) for i in range(min(8, len(lengths)))]
    wavs = [base[i % len(base)][: int(n)] for i, n in enumerate(lengths)]
    starts = np.zeros(len(wavs), np.int64)
    starts[1:] = np.cumsum((lengths[:-1] + 3) // 4 * 4)
    total = int(starts[-1] + (lengths[-1] + 3) // 4 * 4)
    host_pcm = torch.zeros(total, dtype=torch.float32).pin_memory()
    for w, s0, n in zip(wavs, starts, lengths):
        host_pcm[s0:s0 + n] = w
    dev_pcm = host_pcm.to(dev)
    rows = int(T.sum())
    out_dev = torch.empty((rows, AD_CONF["llm_dim"]), dtype=torch.bfloat16, device=dev)
    out_host = torch.empty((rows, AD_CONF["llm_dim"]), dtype=torch.bfloat16).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def step_device():
        pipe.embed_device(dev_pcm, starts, lengths, out=out_dev)

    def step_e2e():
        pipe.run_pinned(host_pcm, starts, lengths, out_host, overlap=False if args.no_overlap else None)

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

    for _ in range(max(args.warmup, 3)):
        step_device()
    torch.cuda.synchronize()
    # ---- launches per step + per-kernel-class device time (profiled steps, NOT the timed ones)
    # (single-stream schedule while profiling: event pairs around launches that overlap on two streams would double count)
    tower.plan().set_streams(1)
    _lib.profile_enable(True)
    for _ in range(2):
        flush.zero_()
        step_device()
    prof = _lib.profile_collect()
    _lib.profile_enable(False)
    tower.plan().set_streams(args.streams if args.streams is not None else 2)
    step_device()
    prof = {k: (ms / 2, n // 2) for k, (ms, n) in prof.items()}

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
    if world > 1:
        t = torch.tensor([ms_dev, ms_e2e], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_dev, ms_e2e = t.tolist()

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak_tf, peak_src = (peaks.get("bf16_tflops_sustained"), "measured sustained") if peaks else (1400.0, "fallback sustained")
        gemm_flop, attn_flop = flops_of(T)
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
        gemm_ms, gemm_n = prof["gemm"]
        achieved = gemm_flop / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
        line = {
            "metric": METRIC, "value": world * audio_s / (ms_dev * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": args.workload, "utterances_per_gpu": int(len(lengths)), "audio_s_per_gpu": audio_s,
                       "tokens_per_gpu": rows, "model": "SenseVoice SAN-M 50+20 blocks d512 ff2048 + adaptor 2x(d1024,8h), random init",
                       "frontend": "povey, 80 mel, LFR 7/6, CMVN", "l2": "256 MiB flush before every timed step; weights 467 MB > L2",
                       "sharding": "by utterance, no collective",
                       "host_binding": (f"rank bound to {len(numa_cores)} GPU-local cores" if numa_cores else "none"),
                       "slab_tokens": args.chunk_tokens if args.chunk_tokens is not None else "off (library default)",
                       "streams": args.streams if args.streams is not None else 2,
                       "e2e_schedule": "single pass" if (args.no_overlap or rows < 8192 or len(lengths) < 8) else
                       "two half-batches on two streams: H2D of half 2 and D2H of half 1 under compute"},
            "e2e": {"value": world * audio_s / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(host_pcm.numel() * 4), "d2h_bytes_per_step": int(out_host.numel() * 2)},
            "gpu_launches": int(sum(n for _, n in prof.values())),
            "roofline": {"bound": "tensor", "kernel": "gemm_bf16_kernel (tcgen05, all 70x4+10 GEMMs of a step)", "achieved": achieved,
                         "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved / peak_tf if peak_tf else None, "traffic": traffic_gemm,
                         "peak_source": peak_src, "launches": gemm_n, "ms_per_step": gemm_ms,
                         # the profiled step brackets every launch with an event pair, which serialises the launches
                         # (no programmatic-dependent-launch overlap); the same class time scaled to the timed step:
                         "achieved_share_of_timed_step": (gemm_flop / (gemm_ms * ms_dev / max(sum(ms for ms, _ in prof.values()), 1e-9) * 1e-3) / 1e12
                                                          if gemm_ms > 0 else 0.0)},
            "kernel_ms_per_step": {k: round(ms, 4) for k, (ms, n) in prof.items()},
            "kernel_launches_per_step": {k: n for k, (ms, n) in prof.items()},
            "clocks": clocks,
        }
        fe_ms = prof["frontend"][0]
        if fe_ms > 0:
            fe_bytes = float(4 * lengths.sum() + 4 * 560 * T.sum())
            hbm = peaks.get("hbm_gbs", 6650.0)
            line["roofline_frontend"] = {"bound": "hbm", "kernel": "fbank_kernel<0>", "achieved": fe_bytes / (fe_ms * 1e-3) / 1e9,
                                         "peak": hbm, "unit": "GB/s", "frac": fe_bytes / (fe_ms * 1e-3) / 1e9 / hbm, "traffic": traffic_fe}
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(lengths, cmvn)
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg5", choices=["cfg5", "cfg2", "cfg3"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-overlap", action="store_true", help="e2e leg: one pass instead of the two-half overlapped schedule")
    ap.add_argument("--chunk-tokens", type=int, default=None, help="slab size of the layer schedule (0 = whole batch per layer)")
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
