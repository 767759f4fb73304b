"""Is a small batch latency-bound enough that k independent graph replays (k sub-batches on k streams, k twin plans) beat one replay
of the whole batch?  Full 50+20 tower, 16 x 10 s (cfg2) and 45 x 10 s (one hour over 8 GPUs), encoder + adaptor only.
usage: python profiles/scripts/small_batch_split.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path[:0] = [ROOT, os.path.join(ROOT, "fun-asr-vllm_b200"), os.path.join(ROOT, "tests")]
import funasr_b200 as fb  # noqa: E402
from oracle import encoder as oe  # noqa: E402

conf = dict(oe.NANO_ENCODER_CONF)
enc = fb.SenseVoiceEncoderSmall(560, **conf)
ad = fb.Transformer(**oe.NANO_ADAPTOR_CONF)
oe.seeded_init_(enc, seed=1), oe.seeded_init_(ad, seed=2)
enc.cuda(), ad.cuda()
tower = fb.AudioTower(enc, ad, max_tokens=8192)
base = tower.plan()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for n_utts in (16, 45):
    lens = np.full(n_utts, 167, np.int32)
    M = int(lens.sum())
    feats = torch.randn(M, 560, device="cuda")
    out = torch.empty(M, 1024, dtype=torch.bfloat16, device="cuda")
    ref = None
    for k in (1, 2, 3, 4):
        plans = [base] + [base.twin() for _ in range(k - 1)]
        streams = [torch.cuda.Stream() for _ in range(k)]
        cuts = np.linspace(0, n_utts, k + 1).astype(int)

        def step():
            cur = torch.cuda.current_stream()
            for i in range(k):
                a, b = cuts[i], cuts[i + 1]
                streams[i].wait_stream(cur)
                with torch.cuda.stream(streams[i]):
                    plans[i].encode_adapt(feats[a * 167:b * 167], 0, lens[a:b], out[a * 167:b * 167], 0)
            for s in streams:
                cur.wait_stream(s)

        for _ in range(3):
            step()
        torch.cuda.synchronize()
        ts = []
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); step(); e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        if ref is None:
            ref = out.clone()
        same = bool(torch.equal(out, ref))
        print(f"{n_utts} utterances ({M} rows) in {k} sub-batch(es): best {min(ts):.3f} ms, median {sorted(ts)[5]:.3f} ms, identical to one batch: {same}", flush=True)
