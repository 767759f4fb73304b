"""Kernel probe used for the ncu captures under profiles/: runs each hot kernel alone at cfg5 size (1 h of audio,
60 120 tokens) with CUDA-event timing.  `python profiles/scripts/probe_kernels.py [frontend|gemm|attn|fsmn|ln|all]`"""
import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fun-asr-vllm_b200"), os.path.join(ROOT, "tests")]
import funasr_b200 as fb  # noqa: E402
import gpu_util as U  # noqa: E402

which = sys.argv[1] if len(sys.argv) > 1 else "all"
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
M = 60120
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, reps=reps):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts), sum(ts) / len(ts)


if which in ("frontend", "all"):
    fe = fb.WavFrontend(window="povey", lfr_m=7, lfr_n=6, dither=0.0, cmvn=torch.zeros(2, 560) + 0.5, device="cuda")
    n = np.full(360, 160000, np.int64)
    starts = np.arange(360, dtype=np.int64) * 160000
    wav = (torch.rand(360 * 160000, device="cuda") - 0.5)
    out = torch.empty((360 * 167, 560), device="cuda")
    best, avg = timeit(lambda: fe.forward_packed(wav, starts, n, out=out))
    byt = 4 * n.sum() + 4 * 560 * 167 * 360
    print(f"frontend 1h: best {best*1e3:.1f} us avg {avg*1e3:.1f} us -> {byt/best/1e6:.0f} GB/s algorithmic")

if which in ("gemm", "all"):
    for name, N, K, relu, f32, nres in [("qkv", 1536, 512, 0, 0, 0), ("out", 512, 512, 0, 1, 1), ("ffn1", 2048, 512, 1, 0, 0),
                                         ("ffn2", 512, 2048, 0, 1, 1), ("qkv0", 1536, 560, 0, 0, 0)]:
        a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
        w = (torch.randn(N, K, device="cuda") / math.sqrt(K)).to(torch.bfloat16)
        bias = torch.randn(N, device="cuda")
        r1 = torch.randn(M, N, device="cuda") if nres >= 1 else None
        r2 = torch.randn(M, N, device="cuda") if nres >= 2 else None
        out = None
        if nres >= 1:  # the plan's form: the last residual IS the output (in-place accumulate of the residual stream)
            out = r2 if nres >= 2 else r1
        best, avg = timeit(lambda: U.op_gemm(a, w, bias, r1, r2, relu=bool(relu), out_f32=bool(f32), out=out))
        print(f"gemm {name:5s} M={M} N={N} K={K}: best {best*1e3:.1f} us avg {avg*1e3:.1f} us -> {2*M*N*K/best/1e9:.0f} TFLOP/s")

if which in ("gemmln", "all"):
    # folded-LayerNorm epilogues: producer (x += ..., bf16 copy, row partials) and consumer (normalise in the epilogue)
    for name, K in [("out+ln", 512), ("ffn2+ln", 2048)]:
        # the out-projection reads [ctx | mem] rows: A = left half, bf16 residual = right half (row stride 2 K)
        cm = torch.randn(M, 2 * K if K == 512 else K, device="cuda").to(torch.bfloat16)
        a = cm[:, :K]
        mem = cm[:, K:] if K == 512 else None
        w = (torch.randn(512, K, device="cuda") / math.sqrt(K)).to(torch.bfloat16)
        bias, x = torch.randn(512, device="cuda"), torch.randn(M, 512, device="cuda")
        xb_, st_ = torch.empty(M, 512, device="cuda", dtype=torch.bfloat16), torch.zeros(M, 16, 2, device="cuda")
        best, avg = timeit(lambda: U.op_gemm_ln_producer(a, w, bias, x, xb=xb_, stats=st_, res_bf16=mem))
        print(f"gemm {name:8s} M={M} N=512 K={K}: best {best*1e3:.1f} us avg {avg*1e3:.1f} us (producer{', bf16 residual' if mem is not None else ''})")
        if mem is not None:
            best, avg = timeit(lambda: U.op_gemm_ln_producer(a, w, bias, x, xb=xb_, stats=st_))
            print(f"gemm {name:8s} M={M} N=512 K={K}: best {best*1e3:.1f} us avg {avg*1e3:.1f} us (producer, without the bf16 residual)")
        r1 = torch.randn(M, 512, device="cuda")
        best, avg = timeit(lambda: U.op_gemm(a, w, bias, r1, None, out_f32=True, out=x))
        print(f"gemm {name[:-3]:8s} M={M} N=512 K={K}: best {best*1e3:.1f} us (staged residual, no LN outputs)")
        best, avg = timeit(lambda: U.op_gemm(a, w, bias, x, None, out_f32=True, out=x))
        print(f"gemm {name[:-3]:8s} M={M} N=512 K={K}: best {best*1e3:.1f} us (accumulate)")
    _, xb, stats, n = U.op_gemm_ln_producer(a, w, bias, x)
    for name, N, relu in [("ln+qkv", 1536, False), ("ln+ffn1", 2048, True)]:
        wf = (torch.randn(N, 512, device="cuda") / math.sqrt(512)).to(torch.bfloat16)
        bb, cs = torch.randn(N, device="cuda"), wf.float().sum(dim=1)
        out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        best, avg = timeit(lambda: U.op_gemm_ln_consumer(xb, wf, bb, cs, stats, n, relu=relu, out=out))
        print(f"gemm {name:8s} M={M} N={N} K=512: best {best*1e3:.1f} us avg {avg*1e3:.1f} us -> {2*M*N*512/best/1e9:.0f} TFLOP/s (consumer)")

if which in ("attn", "all"):
    for H, lens in [(4, [167] * 360), (8, [167] * 360), (4, [500] * 120)]:
        cu = U.cu_of(lens)
        qkv = torch.randn(int(cu[-1]), 3 * H * 128, device="cuda").to(torch.bfloat16)
        best, avg = timeit(lambda: U.op_attention(qkv, cu, H))
        fl = 4 * H * 128 * float((np.array(lens, dtype=np.float64) ** 2).sum())
        print(f"attention H={H} T={lens[0]} x{len(lens)}: best {best*1e3:.1f} us -> {fl/best/1e9:.0f} TFLOP/s")

if which in ("attnfsmn", "all"):
    # the plan's form: attention + FSMN memory in one kernel, the memory added into the f32 residual stream
    lens = [167] * 360
    cu = U.cu_of(lens)
    qkv = torch.randn(M, 1536, device="cuda").to(torch.bfloat16)
    wt = (torch.randn(11, 512, device="cuda") / 3.3)
    x = torch.zeros(M, 512, device="cuda")
    best, avg = timeit(lambda: U.op_attention_fsmn(qkv, wt, cu, 4, 5, acc=x))
    print(f"attention+fsmn (memory into x) H=4 T=167 x360: best {best*1e3:.1f} us avg {avg*1e3:.1f} us -> {(M*1536*2 + M*512*2 + M*512*8)/best/1e6:.0f} GB/s algorithmic")

if which in ("fsmn", "all"):
    cu = U.cu_of([167] * 360)
    qkv = torch.randn(M, 1536, device="cuda").to(torch.bfloat16)
    w = torch.randn(11, 512, device="cuda")
    cm, cu_d = torch.zeros(M, 1024, device="cuda", dtype=torch.bfloat16), U.dev(cu)
    best, avg = timeit(lambda: U.op_fsmn_bf16(qkv, 1024, w, cu, 512, 11, 5, out=cm, cu_d=cu_d))
    print(f"fsmn: best {best*1e3:.1f} us -> {(M*512*4)/best/1e6:.0f} GB/s algorithmic (bf16 in, bf16 out beside ctx)")

if which in ("ln", "all"):
    x = torch.randn(M, 512, device="cuda")
    g, b = torch.ones(512, device="cuda"), torch.zeros(512, device="cuda")
    best, avg = timeit(lambda: U.op_layernorm(x, g, b, True, False))
    print(f"layernorm: best {best*1e3:.1f} us -> {(M*512*6)/best/1e6:.0f} GB/s algorithmic (f32 in, bf16 out)")

if which in ("resample", "all"):
    for sr in (44100, 48000, 8000):
        rs = fb.Resample(sr, 16000)
        n = np.full(360, 10 * sr, np.int64)
        starts = np.arange(360, dtype=np.int64) * (10 * sr)
        wav = torch.rand(int(n.sum()), device="cuda") - 0.5
        on = rs.out_lengths(n)
        out = torch.zeros(int(((on + 3) // 4 * 4).sum()), device="cuda")
        best, avg = timeit(lambda: rs.forward_packed(wav, starts, n, out=out))
        byt = 4 * (n.sum() + on.sum())
        print(f"resample {sr}->16000 1h ({rs.orig}/{rs.new}, {2*rs.width+rs.orig} taps): best {best*1e3:.1f} us -> {byt/best/1e6:.0f} GB/s algorithmic, "
              f"{2*on.sum()*(2*rs.width+rs.orig)/best/1e9:.1f} TFLOP/s fp32")
