"""Audio ingest: the resampler of /root/reference/infer.py:328-329.  CPU part: the oracle and the
product's filter table against outputs of the real torchaudio transform (tests/golden/resample_golden.npz, generator
tests/golden/make_golden_resample.py).  GPU part (-m gpu): fa_resample_sinc through the `Resample` drop-in and the packed
form.  Tolerance: filter tables bit-exact; samples within 2e-6 absolute for inputs in [-0.5, 0.5) (fp32 sums of up to 475
products in a different order than conv1d's)."""
import os

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "resample_golden.npz")
RATES = (8000, 48000, 44100, 22050, 24000, 32000, 11025)
TOL = 2e-6


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


@pytest.mark.parametrize("sr", RATES)
def test_oracle_matches_torchaudio(golden, sr):
    from oracle import ingest as oi

    k, width, _, _ = oi.sinc_resample_kernel(sr, 16000)
    assert width == int(golden[f"width_{sr}"]) and np.array_equal(k, golden[f"kernel_{sr}"])
    y = oi.resample(golden[f"in_{sr}"], sr, 16000)
    assert y.shape == golden[f"out_{sr}"].shape
    assert np.abs(y - golden[f"out_{sr}"]).max() < TOL


@pytest.mark.parametrize("sr", RATES)
def test_product_filter_table_is_torchaudios(golden, sr):
    from funasr_b200.ingest import Resample

    r = Resample(sr, 16000)
    assert r.width == int(golden[f"width_{sr}"]) and np.array_equal(r._table, golden[f"kernel_{sr}"])
    assert r.out_lengths([len(golden[f"in_{sr}"])])[0] == len(golden[f"out_{sr}"])


def test_same_rate_is_identity_and_bad_arguments_raise():
    from funasr_b200.ingest import Resample

    x = torch.arange(10, dtype=torch.float32)
    assert Resample(16000, 16000)(x) is x
    with pytest.raises(ValueError):
        Resample(8000, 16000, resampling_method="sinc_interp_kaiser")
    with pytest.raises(TypeError):
        Resample(8000, 16000)(torch.arange(10))


@pytest.mark.gpu
@pytest.mark.parametrize("sr", RATES)
def test_gpu_resample_vs_torchaudio_golden(golden, sr):
    from funasr_b200.ingest import Resample

    y = Resample(sr, 16000)(torch.from_numpy(golden[f"in_{sr}"]))
    assert y.is_cuda and y.shape == golden[f"out_{sr}"].shape
    assert np.abs(y.cpu().numpy() - golden[f"out_{sr}"]).max() < TOL


@pytest.mark.gpu
@pytest.mark.parametrize("sr", [44100, 48000, 8000])
def test_gpu_resample_packed_ragged_vs_oracle(sr):
    """Ragged packed batch (lengths from 1 sample to several tiles), then straight into the fbank frontend."""
    import funasr_b200 as fb
    from oracle import ingest as oi

    g = torch.Generator().manual_seed(sr)
    lens = np.array([1, 7, sr // 100, sr // 3 + 5, 2 * sr + 11, 3000, sr], np.int64)
    starts = np.zeros(len(lens), np.int64)
    starts[1:] = np.cumsum((lens[:-1] + 3) // 4 * 4)
    total = int(starts[-1] + (lens[-1] + 3) // 4 * 4)
    host = torch.zeros(total)
    for s0, n in zip(starts, lens):
        host[s0:s0 + n] = 0.5 * (2 * torch.rand(int(n), generator=g) - 1)
    rs = fb.Resample(sr, 16000)
    out, ostarts, on = rs.forward_packed(host.cuda(), starts, lens)
    out = out.cpu().numpy()
    for b, (s0, n) in enumerate(zip(starts, lens)):
        want = oi.resample(host[s0:s0 + n].numpy(), sr, 16000)
        assert on[b] == len(want)
        assert np.abs(out[ostarts[b]:ostarts[b] + on[b]] - want).max() < TOL, b
        pad = out[ostarts[b] + on[b]:(ostarts[b + 1] if b + 1 < len(lens) else len(out))]
        assert not pad.any()  # alignment gaps stay zero
    # batched [B, L] form == row by row
    x = host[starts[4]:starts[4] + 2 * sr].reshape(2, sr)
    y = rs(x)
    assert y.shape == (2, int(rs.out_lengths([sr])[0]))
    for r in range(2):
        assert np.abs(y[r].cpu().numpy() - oi.resample(x[r].numpy(), sr, 16000)).max() < TOL
    # the 16 kHz packed output feeds the frontend as is (utterances long enough for one frame)
    keep = on >= 400
    fe = fb.WavFrontend(fs=16000, window="povey", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                        cmvn=None, device="cuda")
    feats, cu = fe.forward_packed(torch.from_numpy(out).cuda(), ostarts[keep], on[keep])
    assert feats.shape[0] == cu[-1] and torch.isfinite(feats).all()
