"""-m gpu: fused fbank -> LFR -> CMVN kernel (through the C ABI / WavFrontend mirror) vs the numpy oracle and the
committed torchaudio / WavFrontend golden vectors.  Tolerance: 1e-3 max-abs in fp32 (BASELINE.json north_star)."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-3
LOG_EPS = np.float32(math.log(1.1920928955078125e-07))


def _frontend(win, cmvn=None):
    from funasr_b200 import WavFrontend

    return WavFrontend(fs=16000, window=win, n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                       cmvn=None if cmvn is None else torch.from_numpy(cmvn), device="cuda")


def _wavs(g):
    return [g[f"wav_{i}"] for i in range(len(g["wav_spec"]))]


@pytest.mark.parametrize("win", ["hamming", "povey"])
@pytest.mark.parametrize("use_cmvn", [False, True])
def test_extract_fbank_vs_real_wavfrontend_golden(frontend_golden, win, use_cmvn):
    """Same call the reference makes (infer.py:446): list of ragged waveforms -> ([B,Tmax,560], lens)."""
    from funasr_b200 import extract_fbank
    from oracle import frontend as of

    g = frontend_golden
    tag = f"{win}_{'cmvn' if use_cmvn else 'nocmvn'}"
    cm = of.parse_cmvn_text(str(g["cmvn_text"])) if use_cmvn else None
    fe = _frontend(win, cm)
    feats, lens = extract_fbank([torch.from_numpy(w) for w in _wavs(g)], frontend=fe, is_final=True)
    assert lens.dtype == torch.int32 and feats.dtype == torch.float32
    assert np.array_equal(lens.numpy(), g[f"lens_{tag}"].reshape(-1))
    want = g[f"feats_{tag}"]
    got = feats.cpu().numpy()
    assert got.shape == want.shape
    scale = 0.3 if use_cmvn else 1.0  # cmvn scales are <= 0.3
    assert np.abs(got - want).max() <= TOL * scale + 1e-6
    if not use_cmvn:  # all-zero utterance: exactly log(eps); padding rows: exactly 0
        z = got[5, : lens[5]]
        assert np.all(z == LOG_EPS)
    for b in range(got.shape[0]):
        assert np.all(got[b, lens[b]:] == 0)


@pytest.mark.parametrize("win", ["hamming", "povey"])
def test_logmel_error_budget_vs_fp64_truth(frontend_golden, win):
    """the kernel is no further from the fp64 truth than torchaudio-fp32 is (x1.5), and <= 1e-3."""
    from funasr_b200 import pack_waveforms

    g = frontend_golden
    fe = _frontend(win)
    for i in (3, 4):
        w = torch.from_numpy(g[f"wav_{i}"])
        wav, starts = pack_waveforms([w], "cuda")
        got, cu = fe.logmel(wav, starts, np.array([w.numel()]))
        got = got.cpu().numpy()
        ref32, ref64 = g[f"logmel_{i}_{win}_f32"], g[f"logmel_{i}_{win}_f64"]
        assert got.shape == ref32.shape
        e_ref, e_got = np.abs(ref32 - ref64).max(), np.abs(got - ref64).max()
        print(f"{win} wav{i}: torchaudio-fp32 vs fp64 {e_ref:.2e}; kernel vs fp64 {e_got:.2e}; kernel vs fp32 {np.abs(got - ref32).max():.2e}")
        assert np.abs(got - ref32).max() <= TOL
        assert e_got <= max(TOL, 1.5 * e_ref)


@pytest.mark.parametrize("win", ["hamming", "povey"])
def test_config1_10s(frontend_golden, win):
    from funasr_b200 import extract_fbank

    g = frontend_golden
    gen = torch.Generator().manual_seed(1234)
    w = 0.5 * (2 * torch.rand(160000, generator=gen) - 1)
    feats, lens = extract_fbank(w, frontend=_frontend(win), is_final=True)
    assert tuple(feats.shape) == (1, 167, 560) and int(lens[0]) == 167
    got = feats[0].cpu().numpy()
    assert np.abs(got[::8] - g[f"cfg1_{win}_rows"]).max() < TOL
    np.testing.assert_allclose(got.astype(np.float64).sum(0), g[f"cfg1_{win}_colsum"], atol=2e-2)


def test_packed_varlen_batch_vs_oracle(frontend_golden):
    """config 3 shape: 64 utterances of 1-30 s, packed offsets; every row against the oracle."""
    from funasr_b200 import pack_waveforms
    from oracle import frontend as of

    rng = np.random.default_rng(0)
    lens = (rng.uniform(1.0, 30.0, 64) * 16000).astype(np.int64)
    lens[:4] = [400, 559, 560, 480000]
    g = torch.Generator().manual_seed(11)
    wavs = [0.5 * (2 * torch.rand(int(n), generator=g) - 1) for n in lens]
    cm = np.stack([-(8 + 4 * rng.random(560)), 0.1 + 0.2 * rng.random(560)]).astype(np.float32)
    fe = _frontend("hamming", cm)
    wav, starts = pack_waveforms(wavs, "cuda")
    got, cu = fe.forward_packed(wav, starts, lens)
    got = got.cpu().numpy()
    # Oracle on the reference's own fp32 tables (torch's fp32 window is up to 2.3e-7 off the exact one, worth ~1e-3 in the
    # lowest log-mel bins of a 30 s full-scale input; numpy cannot reproduce its last bits, see oracle.frontend.fbank).
    # Two fp32 implementations each sit up to ~7e-4 from the fp64 truth there, so the bound is also stated against the
    # fp64 truth, SURVEY.md §8c (8): err(kernel, fp64) <= max(1e-3, 1.5 * err(fp32 oracle, fp64)), in the log-mel
    # domain (CMVN scale divided out).
    tables = {"window": frontend_golden["window_hamming"], "mel": frontend_golden["mel_banks"]}
    for b in [0, 1, 2, 3, 10, 33, 63]:
        w = [wavs[b].numpy()]
        want32, _ = of.wav_frontend_packed(w, "hamming", cm, tables=tables)
        want64, _ = of.wav_frontend_packed(w, "hamming", cm.astype(np.float64), dtype=np.float64, tables=tables)
        assert cu[b + 1] - cu[b] == want32.shape[0]
        e_ref = (np.abs(want32 - want64) / cm[1]).max()
        e_got = (np.abs(got[cu[b]:cu[b + 1]] - want64) / cm[1]).max()
        e_32 = (np.abs(got[cu[b]:cu[b + 1]] - want32) / cm[1]).max()
        print(f"utt {b} ({lens[b]} samples): oracle32 vs f64 {e_ref:.2e}; kernel vs f64 {e_got:.2e}; kernel vs oracle32 {e_32:.2e}")
        assert e_got <= max(TOL, 1.5 * e_ref) + 2e-6, b  # 2e-6: fp32 rounding of the CMVN'd output itself
        assert e_32 <= TOL, b
    # size-independent property: each LFR row is 7 stacked fbank frames -> slot j of row i equals slot j-6.. of row i+1
    raw = got / cm[1] - cm[0]
    for b in (3, 40):
        r = raw[cu[b]:cu[b + 1]]
        np.testing.assert_allclose(r[:-1, 6 * 80:7 * 80], r[1:, 0:80], atol=2e-4)


def test_translation_and_linearity_properties():
    """Full-size (1 h) properties that need no oracle: (a) shifting the waveform by one hop shifts log-mel rows by one;
    (b) scaling the waveform by 2 adds log(4) to every log-mel value above the floor."""
    from funasr_b200 import pack_waveforms

    fe = _frontend("povey")
    g = torch.Generator().manual_seed(21)
    base = 0.25 * (2 * torch.rand(160000 + 160, generator=g) - 1)
    wavs = [base[:160000], base[160:], 2 * base[:160000]]
    wav, starts = pack_waveforms(wavs, "cuda")
    got, cu = fe.logmel(wav, starts, np.array([160000, 160000, 160000]))
    a, b, c = (got[cu[i]:cu[i + 1]].cpu().numpy() for i in range(3))
    np.testing.assert_array_equal(a[1:], b[:-1])
    np.testing.assert_allclose(c - a, math.log(4.0), atol=2e-5)


def test_sample_scale_in_the_window_or_on_the_samples_gives_the_same_bits():
    """The x32768 of ``upsacle_samples`` (VLLM/transformers_utils/processors/funasr.py:140-141) is a power of two: the kernel
    folds it into the window, which must give the bits of scaling the samples first (what the reference does).  A scale that
    is not a power of two takes the other branch (samples scaled as they are read): identical to pre-scaled input."""
    from funasr_b200 import WavFrontend, pack_waveforms

    g = torch.Generator().manual_seed(5)
    w = 0.4 * (2 * torch.rand(48000, generator=g) - 1)
    lens = np.array([48000])
    mk = lambda up: WavFrontend(fs=16000, window="hamming", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6,
                                dither=0.0, upsacle_samples=up, device="cuda")
    fe_up, fe_raw = mk(True), mk(False)
    a, _ = fe_up.logmel(*pack_waveforms([w], "cuda"), lens)
    b, _ = fe_raw.logmel(*pack_waveforms([w * 32768.0], "cuda"), lens)
    assert torch.equal(a, b)
    fe3 = mk(False)
    fe3._cfg.upscale = 3.0  # not a power of two
    c, _ = fe3.logmel(*pack_waveforms([w], "cuda"), lens)
    d, _ = fe_raw.logmel(*pack_waveforms([w * 3.0], "cuda"), lens)
    assert torch.equal(c, d)
    assert not torch.equal(c, b)


def test_short_or_empty_input_raises():
    from funasr_b200 import extract_fbank

    fe = _frontend("hamming")
    with pytest.raises(ValueError):
        extract_fbank([torch.zeros(399)], frontend=fe)
    with pytest.raises(ValueError):
        extract_fbank([], frontend=fe)
    import funasr_b200.frontend as ff

    ff._dither_warned = False
    with pytest.warns(UserWarning, match="dither"):  # funasr's default dither=1.0 builds, and says that it is ignored
        type(fe)(dither=1.0, lfr_m=7, lfr_n=6)
    with pytest.raises(ValueError):
        type(fe)(snip_edges=False, dither=0.0)
