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


def test_oracle_pcm_to_mono_matches_torch_and_numpy_mean():
    """The collator's cast is torch's (.float(), infer.py:320-324); the mono mix is numpy's mean over channels."""
    from oracle import ingest as oi

    rng = np.random.default_rng(5)
    s16 = rng.integers(-32768, 32767, size=(1001, 2), dtype=np.int16)
    f64 = rng.standard_normal(777) * 0.3
    f32 = (rng.standard_normal((640, 3)) * 0.2).astype(np.float32)
    assert np.array_equal(oi.pcm_to_mono_f32(s16[:, 0]), torch.from_numpy(s16[:, 0].copy()).float().numpy())
    assert np.array_equal(oi.pcm_to_mono_f32(f64), torch.from_numpy(f64).float().numpy())
    for x, sc in ((s16, 1.0 / 32768), (f32, 1.0)):
        want = np.mean(x.astype(np.float32).T * np.float32(sc), axis=0)  # librosa.to_mono on soundfile's [L, C] transposed
        assert np.array_equal(oi.pcm_to_mono_f32(x, sc), want.astype(np.float32))


@pytest.mark.gpu
@pytest.mark.parametrize("dtype,ch,scale", [(np.int16, 1, 1.0), (np.int16, 2, 1.0 / 32768), (np.float64, 1, 1.0), (np.float32, 3, 1.0),
                                            (np.int16, 6, 1.0 / 32768)])
def test_gpu_pcm_to_mono_bit_exact(dtype, ch, scale):
    import funasr_b200 as fb
    from oracle import ingest as oi

    rng = np.random.default_rng(ch * 7 + np.dtype(dtype).itemsize)
    lens = [1, 4097, 160000, 333, 16]
    shape = (lambda n: (n,) if ch == 1 else (n, ch))
    if dtype == np.int16:
        waves = [rng.integers(-32768, 32767, size=shape(n), dtype=np.int16) for n in lens]
    else:
        waves = [(rng.standard_normal(shape(n)) * 0.3).astype(dtype) for n in lens]
    host, fmt, c, starts, n = fb.pack_pcm(waves)
    assert c == ch
    out, ostarts, on = fb.pcm_to_mono(host.cuda(), fmt, c, starts, n, scale)
    out = out.cpu().numpy()
    for w, s0, k in zip(waves, ostarts, on):
        assert np.array_equal(out[s0:s0 + k], oi.pcm_to_mono_f32(w, scale))


@pytest.mark.gpu
@pytest.mark.parametrize("sr,dtype,ch,scale", [(44100, np.int16, 2, 1.0 / 32768), (48000, np.float64, 1, 1.0), (8000, np.int16, 1, 1.0)])
def test_gpu_resample_from_raw_pcm_is_the_two_step_result(sr, dtype, ch, scale):
    """fa_resample_sinc_pcm = fa_resample_sinc(fa_pcm_to_mono_f32(raw)) bit for bit (same staged values, same taps), and
    both follow the oracle (float cast + mono mean, then torchaudio's polyphase sinc)."""
    import funasr_b200 as fb
    from oracle import ingest as oi

    rng = np.random.default_rng(sr)
    lens = [sr // 3, sr * 2 + 17, 5]
    shape = (lambda n: (n,) if ch == 1 else (n, ch))
    if dtype == np.int16:
        waves = [rng.integers(-20000, 20000, size=shape(n), dtype=np.int16) for n in lens]
    else:
        waves = [(rng.standard_normal(shape(n)) * 0.3).astype(dtype) for n in lens]
    host, fmt, c, starts, n = fb.pack_pcm(waves)
    rs = fb.Resample(sr, 16000)
    fused, fs, fn = rs.forward_packed_pcm(host.cuda(), fmt, c, starts, n, scale)
    mono, ms, mn = fb.pcm_to_mono(host.cuda(), fmt, c, starts, n, scale)
    two, ts, tn = rs.forward_packed(mono, ms, mn)
    assert np.array_equal(fs, ts) and np.array_equal(fn, tn)
    fused, two = fused.cpu().numpy(), two.cpu().numpy()
    for w, s0, k in zip(waves, fs, fn):
        assert np.array_equal(fused[s0:s0 + k], two[s0:s0 + k])
        want = oi.resample(oi.pcm_to_mono_f32(w, scale), sr, 16000)
        assert k == len(want)
        tol = 2e-6 * max(1.0, float(np.abs(want).max()))
        assert np.abs(fused[s0:s0 + k] - want).max() <= 20 * tol
