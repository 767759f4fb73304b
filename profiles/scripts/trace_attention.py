"""Per-CTA timeline of the attention kernel (debug hook fa_debug_attention_trace): where a CTA's lifetime goes."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fun-asr-vllm_b200"), os.path.join(ROOT, "tests")]
import gpu_util as U  # noqa: E402
from funasr_b200 import _lib  # noqa: E402

lib = _lib.load()
H, lens = 4, [167] * 360
cu = U.cu_of(lens)
qkv = torch.randn(int(cu[-1]), 3 * H * 128, device="cuda").to(torch.bfloat16)
U.op_attention(qkv, cu, H)
torch.cuda.synchronize()
buf = torch.zeros(512 * 16, dtype=torch.int64, device="cuda")
lib.fa_debug_attention_trace.argtypes = [C.c_void_p]
lib.fa_debug_attention_trace(buf.data_ptr())
U.op_attention(qkv, cu, H)
torch.cuda.synchronize()
lib.fa_debug_attention_trace(None)
t = buf.cpu().numpy().reshape(512, 16)
t0 = t[:, 0].min()
names = ["entry", "setup", "q+k0", "s0", "s1", "s2", "p0", "p1", "p2", "o_full", "epi", "end"]
print("cta  sm  " + " ".join(f"{n:>7s}" for n in names) + "   (ns since first CTA entry; deltas from entry)")
for c in list(range(0, 12)) + list(range(296, 304)):
    r = t[c]
    print(f"{c:3d} {int(r[12]):3d}  {int(r[0]-t0):7d} " + " ".join(f"{int(r[i]-r[0]) if r[i] else -1:7d}" for i in range(1, 12)))
life = (t[:, 11] - t[:, 0])[t[:, 11] > 0]
print("CTA lifetime ns: mean", life.mean(), "median", np.median(life), "p90", np.percentile(life, 90))
for i in range(1, 12):
    ok = (t[:, i] > 0) & (t[:, i - 1] > 0)
    print(f"  {names[i-1]:>7s} -> {names[i]:<7s}: mean {np.mean((t[:, i] - t[:, i - 1])[ok]):8.0f} ns  (n={ok.sum()})")
