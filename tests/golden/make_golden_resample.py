"""Generates tests/golden/resample_golden.npz with the REAL torchaudio transform the reference calls
(/root/reference/infer.py:328-329: ``torchaudio.transforms.Resample(sr, 16000)(audio_tensor)``), run in the build
container (torchaudio 2.11).  Small seeded inputs; the filter tables are stored too so the table arithmetic is pinned bit
for bit.  Usage: python tests/golden/make_golden_resample.py"""
import os

import numpy as np
import torch
import torchaudio
from torchaudio.functional.functional import _get_sinc_resample_kernel
import math

out = {}
g = torch.Generator().manual_seed(2024)
for sr, n in [(8000, 1500), (48000, 4801), (44100, 5000), (22050, 2999), (24000, 3000), (32000, 17), (11025, 2000), (16000, 100)]:
    x = 0.5 * (2 * torch.rand(n, generator=g) - 1)
    y = torchaudio.transforms.Resample(sr, 16000)(x)
    out[f"in_{sr}"] = x.numpy()
    out[f"out_{sr}"] = y.numpy()
    if sr != 16000:
        k, w = _get_sinc_resample_kernel(sr, 16000, math.gcd(sr, 16000))
        out[f"kernel_{sr}"] = k[:, 0, :].numpy()
        out[f"width_{sr}"] = np.int64(w)
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "resample_golden.npz"), **out)
print({k: v.shape for k, v in out.items()})
