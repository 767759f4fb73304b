"""Times the attention kernel alone (fa_op_attention), the fused attention+FSMN kernel (fa_op_attention_fsmn) and the
stand-alone FSMN (fa_op_fsmn_bf16) on the packed shapes of the bench workloads.  CUDA events on the launching stream, 256 MiB
L2 flush before every timed launch.  usage: python profiles/scripts/attn_bench.py [iters] [shape ...]"""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, "tests")
sys.path.insert(0, "fun-asr-vllm_b200")
import gpu_util as U  # noqa: E402
from funasr_b200 import _lib  # noqa: E402


def shapes():
    rng = np.random.default_rng(0)
    return {
        "cfg5": [167] * 360,
        "cfg2": [167] * 16,
        "cfg3": [int(x * 16000 / 160 / 6) + 1 for x in rng.uniform(1.0, 30.0, 64)],
        "long": [3000] * 20,
    }


def time_it(fn, flush, iters):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for _ in range(3):
        fn()
    for a, b in ev:
        flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    t = sorted(a.elapsed_time(b) for a, b in ev)
    return t[len(t) // 2] * 1e3


def main():
    iters = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    want = sys.argv[2:] or list(shapes())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    H, dm = 4, 512
    lib, st = U.lib(), U.stream()
    out = {}
    for name in want:
        lens = shapes()[name]
        cu = U.cu_of(lens)
        M = int(cu[-1])
        g = torch.Generator().manual_seed(1)
        qkv = torch.randn(M, 3 * dm, generator=g).to(torch.bfloat16).cuda()
        w = (torch.randn(11, dm, generator=g) / 3.3).cuda()
        cu_d = U.dev(cu)
        cm = torch.zeros((M, 2 * dm), dtype=torch.bfloat16, device="cuda")
        mx = int(max(lens))

        def attn():
            _lib.check(lib.fa_op_attention(qkv.data_ptr(), 3 * dm, cu_d.data_ptr(), len(lens), mx, H, M, cm.data_ptr(), 2 * dm, st))

        def fused():
            _lib.check(lib.fa_op_attention_fsmn(qkv.data_ptr(), 3 * dm, cu_d.data_ptr(), len(lens), mx, H, M, cm.data_ptr(), 2 * dm,
                                                w.data_ptr(), 11, 5, cm.data_ptr() + 2 * dm, 2 * dm, None, 0, st))

        xacc = torch.zeros((M, dm), dtype=torch.float32, device="cuda")

        def fused_acc():  # the memory added straight into an f32 residual stream (vector red) instead of the bf16 store
            _lib.check(lib.fa_op_attention_fsmn(qkv.data_ptr(), 3 * dm, cu_d.data_ptr(), len(lens), mx, H, M, cm.data_ptr(), 2 * dm,
                                                w.data_ptr(), 11, 5, None, 0, xacc.data_ptr(), dm, st))

        def fsmn():
            _lib.check(lib.fa_op_fsmn_bf16(qkv.data_ptr(), 3 * dm, 2 * dm, w.data_ptr(), cu_d.data_ptr(), len(lens),
                                           cm.data_ptr() + 2 * dm, 2 * dm, M, dm, 11, 5, st))

        bytes_attn = M * 3 * dm * 2 + M * dm * 2
        out[name] = {"rows": M, "utts": len(lens), "attention_us": round(time_it(attn, flush, iters), 1),
                     "fused_us": round(time_it(fused, flush, iters), 1), "fused_acc_us": round(time_it(fused_acc, flush, iters), 1), "fsmn_us": round(time_it(fsmn, flush, iters), 1),
                     "hbm_floor_us": round((bytes_attn + M * dm * 2) / 6458.1e3, 1)}
        print(name, json.dumps(out[name]), flush=True)


if __name__ == "__main__":
    main()
