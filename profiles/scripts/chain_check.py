"""First-contact check of the GEMM chain kernel (csrc/gemm_chain.cu): one encoder forward with the chain on against the same
forward with one launch per GEMM, for growing batch sizes; prints max |diff| and timing per case and stops at the first
mismatch.  Run under `timeout`: a scheduling bug traps after ~3 s per stuck wait.
usage: python profiles/scripts/chain_check.py [blocks tp_blocks]"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fun-asr-vllm_b200"), os.path.join(ROOT, "tests")]
import funasr_b200 as fb  # noqa: E402
from oracle import encoder as oe  # noqa: E402  (weights only: seeded init shared with the tests)

nb, ntp = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (4, 2)
conf = dict(oe.NANO_ENCODER_CONF, num_blocks=nb, tp_blocks=ntp)
o_enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **conf), seed=4321)
o_ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)
enc, ad = fb.SenseVoiceEncoderSmall(560, **conf), fb.Transformer(**oe.NANO_ADAPTOR_CONF)
enc.load_state_dict(o_enc.state_dict())
ad.load_state_dict(o_ad.state_dict())
enc.cuda(), ad.cuda()
tower = fb.AudioTower(enc, ad, max_tokens=61000)
cases = [([50], 1, 0), ([167], 1, 0), ([300], 1, 0), ([167] * 16, 1, 0), ([167] * 16, 1, 8192), ([167] * 40 + [500, 33, 257], 1, 0),
         ([167] * 40 + [500, 33, 257], 2, 0), ([167] * 360, 1, 0), ([167] * 360, 2, 0)]
g = torch.Generator().manual_seed(7)
for lens, streams, graph_rows in cases:
    M = sum(lens)
    feats = torch.randn(M, 560, generator=g).cuda()
    il = torch.tensor(lens, dtype=torch.int32)
    plan = tower.plan()
    plan.set_graph_rows(graph_rows)
    plan.set_streams(streams)
    plan.set_chain(False)
    want, _ = tower.forward_packed(feats, il.numpy(), out_dtype=torch.float32)
    want = want.clone()
    torch.cuda.synchronize()
    plan.set_chain(True)
    t0 = time.time()
    got, _ = tower.forward_packed(feats, il.numpy(), out_dtype=torch.float32)
    torch.cuda.synchronize()
    dt = time.time() - t0
    diff = float((got - want).abs().max())
    ts = []
    for on in (False, True):
        plan.set_chain(on)
        for _ in range(2):
            tower.forward_packed(feats, il.numpy(), out_dtype=torch.float32)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            tower.forward_packed(feats, il.numpy(), out_dtype=torch.float32)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / 5)
    print(f"rows {M:6d} utts {len(lens):4d} streams {streams} graph {graph_rows:5d}: max|diff| {diff:.3e} finite {bool(torch.isfinite(got).all())} "
          f"first call {dt*1e3:.1f} ms; per call {ts[0]:.3f} ms unchained, {ts[1]:.3f} ms chained", flush=True)
    if diff != 0.0:
        print("MISMATCH", flush=True)
        sys.exit(1)
print("chain check OK")
