"""Per-tile timeline of one CTA of the GEMM chain kernel (library built with `make variant NAME=ctrace DEFS=-DFA_CHAIN_TRACE`):
runs a 4+2-block tower at a given batch, reads the trace of the last chain launch and prints, per role, where the cycles went.
usage: python profiles/scripts/with_lib.py ctrace profiles/scripts/chain_trace.py [n_utts] [block ...]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fun-asr-vllm_b200"), os.path.join(ROOT, "tests")]
import funasr_b200 as fb  # noqa: E402
from funasr_b200 import _lib  # noqa: E402
from oracle import encoder as oe  # noqa: E402

n_utts = int(sys.argv[1]) if len(sys.argv) > 1 else 360
blocks = [int(a) for a in sys.argv[2:]] or [0, 60]
conf = dict(oe.NANO_ENCODER_CONF, num_blocks=4, tp_blocks=2)
o_enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **conf), seed=4321)
o_ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)
enc, ad = fb.SenseVoiceEncoderSmall(560, **conf), fb.Transformer(**oe.NANO_ADAPTOR_CONF)
enc.load_state_dict(o_enc.state_dict()); ad.load_state_dict(o_ad.state_dict())
enc.cuda(), ad.cuda()
tower = fb.AudioTower(enc, ad, max_tokens=61000)
lens = [167] * n_utts
feats = torch.randn(sum(lens), 560).cuda()
il = np.array(lens, np.int32)
plan = tower.plan()
plan.set_graph_rows(0); plan.set_streams(1); plan.set_chain(True)
lib = _lib.load()
lib.fa_debug_chain_trace.argtypes = [C.c_void_p, C.c_int, C.c_int]
TR_TILES, TR_EV, TR_ROLES = 128, 8, 4
buf = np.zeros(TR_ROLES * TR_TILES * TR_EV, np.int64)
names = ["out", "ffn1", "ffn2", "qkv"]
for blk in blocks:
    lib.fa_debug_chain_trace(buf.ctypes.data, buf.size, blk)  # select the block for the launches that follow
    for _ in range(3):
        tower.forward_packed(feats, il, out_dtype=torch.float32)
    torch.cuda.synchronize()
    lib.fa_debug_chain_trace(buf.ctypes.data, buf.size, blk)
    t = buf.reshape(TR_ROLES, TR_TILES, TR_EV).copy()
    prod, mma, e0, e1 = t
    n = int((prod[:, 3] > 0).sum())
    ids = prod[:n, 0]
    kind = (ids >> 40).astype(int)
    t0 = prod[0, 1]
    span = max(e0[n - 1, 5], e1[n - 1, 5]) - t0
    print(f"=== block {blk}: {n} tiles, span {span} cycles ({span / n:.0f} per tile); tiles by problem {[int((kind == k).sum()) for k in range(4)]}")
    print(f"producer: dep wait {int((prod[:n, 2] - prod[:n, 1]).sum())}  loads (incl. ring back-pressure) {int((prod[:n, 3] - prod[:n, 2]).sum())}")
    print(f"mma: wait acc free {int((mma[:n, 2] - mma[:n, 1]).sum())}  wait first operands {int((mma[:n, 3] - mma[:n, 2]).sum())}  main loops {int((mma[:n, 4] - mma[:n, 3]).sum())}"
          f"  between tiles {int((mma[1:n, 1] - mma[:n - 1, 4]).sum())}")
    for nm, e in (("epi w2", e0), ("epi w6", e1)):
        print(f"{nm}: wait dep {int((e[:n, 2] - e[:n, 1]).sum())}  wait acc full {int((e[:n, 3] - e[:n, 2]).sum())}  work {int((e[:n, 4] - e[:n, 3]).sum())}  publish {int((e[:n, 5] - e[:n, 4]).sum())}")
    for k in range(4):
        m = kind == k
        if m.any():
            print(f"  {names[k]:5s}: main loop {np.mean((mma[:n, 4] - mma[:n, 3])[m]):8.0f}  first operands {np.mean((mma[:n, 3] - mma[:n, 2])[m]):7.0f}  acc-free wait {np.mean((mma[:n, 2] - mma[:n, 1])[m]):7.0f}"
                  f"  epilogue work {np.mean((e0[:n, 4] - e0[:n, 3])[m]):8.0f}  (residual wait {np.mean(e0[:n, 6][m]):6.0f}, store-read wait {np.mean(e0[:n, 7][m]):6.0f})  publish {np.mean((e0[:n, 5] - e0[:n, 4])[m]):7.0f}  producer dep wait {np.mean((prod[:n, 2] - prod[:n, 1])[m]):7.0f}")
    print("first 12 tiles (cycles from start): kind  prod[dep0 dep1 loads]  mma[accwait go first end]  epi[top dep full done pub]")
    for i in range(min(n, 12)):
        print(f"  {names[kind[i]]:5s} rb{int((ids[i] >> 8) & 0xffffffff):4d}  {[int(x - t0) for x in prod[i, 1:4]]}  {[int(x - t0) for x in mma[i, 1:5]]}  {[int(x - t0) for x in e0[i, 1:6]]}")
