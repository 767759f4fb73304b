"""L2-residency probe: each hot kernel at slab size (M rows) launched back to back on the same buffers (no flush, data
stays in the 126 MB L2) vs with a flush before every launch.  `python profiles/scripts/probe_l2.py [M]`"""
import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fun-asr-vllm_b200"), os.path.join(ROOT, "tests")]
import funasr_b200 as fb  # noqa: E402
import gpu_util as U  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 8704
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def hot(fn, n=30):
    fn(); fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3


def cold(fn, n=5):
    ts = []
    for _ in range(n):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return min(ts)


for name, N, K, relu, f32, nres in [("qkv", 1536, 512, 0, 0, 0), ("out", 512, 512, 0, 1, 2), ("ffn1", 2048, 512, 1, 0, 0),
                                     ("ffn2", 512, 2048, 0, 1, 1)]:
    a = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    w = (torch.randn(N, K, device="cuda") / math.sqrt(K)).to(torch.bfloat16)
    bias = torch.randn(N, device="cuda")
    r1 = torch.randn(M, N, device="cuda") if nres >= 1 else None
    r2 = torch.randn(M, N, device="cuda") if nres >= 2 else None
    f = lambda: U.op_gemm(a, w, bias, r1, r2, relu=bool(relu), out_f32=bool(f32))
    h, c = hot(f), cold(f)
    print(f"gemm {name:5s} M={M}: hot {h:.1f} us ({2*M*N*K/h/1e6:.0f} TF)  cold {c:.1f} us ({2*M*N*K/c/1e6:.0f} TF)")
T = 167
nu = M // T
cu = U.cu_of([T] * nu)
Mr = int(cu[-1])
qkv = torch.randn(Mr, 1536, device="cuda").to(torch.bfloat16)
f = lambda: U.op_attention(qkv, cu, 4)
print(f"attention H=4 T=167 x{nu}: hot {hot(f):.1f} us cold {cold(f):.1f} us")
wf = torch.randn(11, 512, device="cuda")
f = lambda: U.op_fsmn(qkv, 1024, wf, cu, 512, 11, 5)
print(f"fsmn: hot {hot(f):.1f} us cold {cold(f):.1f} us")
x = torch.randn(Mr, 512, device="cuda")
g, b = torch.ones(512, device="cuda"), torch.zeros(512, device="cuda")
f = lambda: U.op_layernorm(x, g, b, True, False)
print(f"layernorm: hot {hot(f):.1f} us cold {cold(f):.1f} us")
e = torch.empty(0, device="cuda")
f = lambda: torch.cuda._sleep(1) if False else U.op_layernorm(x[:8], g, b, True, False)
print(f"tiny launch (8-row layernorm) back to back: {hot(f, 200):.2f} us per launch")
