"""Pipeline timeline of the attention(+FSMN) kernel: CTA 0's events with SM clocks, from the profiling build
(`make -C fun-asr-vllm_b200/csrc trace`).  usage: python profiles/scripts/attn_trace.py [fused|plain] [shape] [first_item n_items]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for d in ("tests", "fun-asr-vllm_b200", os.path.join("profiles", "scripts")):
    sys.path.insert(0, os.path.join(ROOT, d))
from funasr_b200 import _lib  # noqa: E402

_lib.LIB_PATH = os.path.join(os.path.dirname(_lib.LIB_PATH), "libfunasr_b200_trace.so")
import gpu_util as U  # noqa: E402
from attn_bench import shapes  # noqa: E402

EV = {1: "Q  issue", 2: "KV issue", 3: "S  issue", 4: "PV issue", 5: "sm wait", 6: "sm go", 7: "P  ready", 8: "epi go",
      9: "fs vfull", 10: "fs rel", 11: "mma item", 12: "mma qfull", 13: "sm pbuf?", 14: "sm pbuf!", 15: "epi end", 16: "mma done"}


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "fused"
    name = sys.argv[2] if len(sys.argv) > 2 else "cfg5"
    first, count = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (4, 2)
    lens = shapes()[name]
    H, dm = 4, 512
    cu = U.cu_of(lens)
    M = int(cu[-1])
    g = torch.Generator().manual_seed(1)
    qkv = torch.randn(M, 3 * dm, generator=g).to(torch.bfloat16).cuda()
    w = (torch.randn(11, dm, generator=g) / 3.3).cuda()
    cu_d = U.dev(cu)
    cm = torch.zeros((M, 2 * dm), dtype=torch.bfloat16, device="cuda")
    xacc = torch.zeros((M, dm), dtype=torch.float32, device="cuda")
    lib, st = U.lib(), U.stream()
    lib.fa_debug_attention_trace.restype = C.c_int
    lib.fa_debug_attention_trace.argtypes = [C.c_void_p, C.c_int]
    buf = np.zeros(1 << 16, np.uint64)

    def run():
        if mode == "fused":
            _lib.check(lib.fa_op_attention_fsmn(qkv.data_ptr(), 3 * dm, cu_d.data_ptr(), len(lens), max(lens), H, M, cm.data_ptr(),
                                                2 * dm, w.data_ptr(), 11, 5, None, 0, xacc.data_ptr(), dm, st))
        else:
            _lib.check(lib.fa_op_attention(qkv.data_ptr(), 3 * dm, cu_d.data_ptr(), len(lens), max(lens), H, M, cm.data_ptr(), 2 * dm, st))

    for _ in range(3):
        run()
    lib.fa_debug_attention_trace(buf.ctypes.data, len(buf))
    run()
    n = lib.fa_debug_attention_trace(buf.ctypes.data, len(buf))
    ev = [(int(x) >> 24, (int(x) >> 16) & 0xff, (int(x) >> 8) & 0xff, int(x) & 0xff) for x in buf[:n] if x != 0]
    n = len(ev)
    ev.sort()
    t0 = ev[0][0]
    print(f"{mode} {name}: {n} events, CTA 0 span {ev[-1][0] - t0} cycles")
    # per-item period from the MMA thread's item starts
    starts = [t for t, e, a, b in ev if e == 11]
    if len(starts) > 2:
        d = np.diff(starts)
        print("item periods (cycles):", " ".join(str(int(x)) for x in d))
    nb = (max(lens) + 63) // 64
    lo, hi = first, first + count
    for t, e, a, b in ev:
        # a = item for producer/MMA events, block counter c for the softmax, global block for the FSMN
        item = a if e in (1, 2, 3, 4, 11, 12, 16) else a // nb
        if lo <= item < hi:
            print(f"{t - t0:9d}  {EV.get(e, e):9s} a={a:3d} b={b:3d}")


if __name__ == "__main__":
    main()
