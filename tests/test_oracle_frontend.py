"""Pins oracle/frontend.py to the real torchaudio / WavFrontend outputs in tests/golden (CPU)."""
import math

import numpy as np
import pytest

from oracle import frontend as of

LOG_EPS = math.log(1.1920928955078125e-07)


def _wavs(g):
    return [g[f"wav_{i}"] for i in range(len(g["wav_spec"]))]


def test_frame_count_known_answers():
    # SURVEY.md §8c (2)
    for n, m, T in [(16000, 98, 17), (160000, 998, 167), (160007, 998, 167), (480000, 2998, 500), (400, 1, 1)]:
        assert of.num_frames(n) == m
        assert of.num_lfr_frames(m) == T
    assert of.num_frames(399) == 0


def test_fake_token_len():
    # /root/reference/model.py:391-396
    assert of.fake_token_len(167, False) == 167
    assert of.fake_token_len(167, True) == 21
    assert of.fake_token_len(17, True) == 3
    assert of.fake_token_len(500, True) == 63


def test_tables_match_torchaudio(frontend_golden):
    g = frontend_golden
    mb = of.mel_banks()
    assert mb.shape == (80, 257)
    assert np.all(mb[:, 256] == 0) and np.all(mb[:, 0] == 0)
    assert (mb > 0).sum() == 501
    assert ((g["mel_banks"] > 0) == (mb[:, :256] > 0)).all()
    np.testing.assert_allclose(mb[:, :256], g["mel_banks"], atol=3e-5)
    nz = (mb > 0).sum(1)
    assert nz.min() == 1 and nz.max() == 16
    for w in ("hamming", "povey"):
        np.testing.assert_allclose(of.window_function(w), g[f"window_{w}"], atol=5e-7)


def test_lfr_closed_form_vs_reference_apply_lfr(frontend_golden):
    g = frontend_golden
    for m in g["lfr_ms"]:
        m = int(m)
        mat = np.arange(m, dtype=np.float32)[:, None]
        got = of.apply_lfr(mat, 7, 6).astype(np.int32)
        assert got.shape == g[f"lfr_{m}"].shape
        assert np.array_equal(got, g[f"lfr_{m}"]), m
        assert np.array_equal(of.apply_lfr_strided(mat, 7, 6).astype(np.int32), got), m


def test_cmvn_known_answer_and_parser(frontend_golden):
    x = np.ones((3, 560), np.float32)
    c = np.stack([np.full(560, -2.0, np.float32), np.full(560, 0.5, np.float32)])
    assert np.all(of.apply_cmvn(x, c) == -0.5)
    cm = of.parse_cmvn_text(str(frontend_golden["cmvn_text"]))
    assert cm.shape == (2, 560) and cm.dtype == np.float32
    assert (cm[0] < 0).all() and (cm[1] > 0).all()


@pytest.mark.parametrize("win", ["hamming", "povey"])
@pytest.mark.parametrize("use_cmvn", [False, True])
def test_wav_frontend_vs_real_wavfrontend(frontend_golden, win, use_cmvn):
    g = frontend_golden
    tag = f"{win}_{'cmvn' if use_cmvn else 'nocmvn'}"
    cm = of.parse_cmvn_text(str(g["cmvn_text"])) if use_cmvn else None
    feats, lens = of.wav_frontend(_wavs(g), win, cm)
    assert np.array_equal(lens, g[f"lens_{tag}"].reshape(-1))
    assert feats.shape == g[f"feats_{tag}"].shape
    # fp32 vs fp32: both sides carry ~5e-4 of rounding noise in the log domain
    scale = 0.3 if use_cmvn else 1.0
    assert np.abs(feats - g[f"feats_{tag}"]).max() <= 1e-3 * scale + 1e-6
    # all-zero utterance (index 5): exactly log(eps) everywhere  
    if not use_cmvn:
        z = feats[5, : lens[5]]
        assert np.all(z == np.float32(LOG_EPS))
        assert np.all(g[f"feats_{tag}"][5, : lens[5]] == np.float32(LOG_EPS))


@pytest.mark.parametrize("win", ["hamming", "povey"])
def test_fp64_truth_and_error_budget(frontend_golden, win):
    g = frontend_golden
    for i in (3, 4):
        w = g[f"wav_{i}"]
        ref32, ref64 = g[f"logmel_{i}_{win}_f32"], g[f"logmel_{i}_{win}_f64"]
        o32, o64 = of.fbank(w, win, np.float32), of.fbank(w, win, np.float64)
        assert np.abs(o64 - ref64).max() < 1e-4  # fp64 restatement == torchaudio fp64 (mel table fp32-built)
        e_ref = np.abs(ref32 - ref64).max()
        e_ora = np.abs(o32 - ref64).max()
        assert e_ref < 1e-3 and e_ora < 1e-3
        assert e_ora <= max(1e-3, 1.5 * e_ref)


def test_windows_are_a_hard_switch(frontend_golden):
    g = frontend_golden
    d = np.abs(g["logmel_3_hamming_f32"] - g["logmel_3_povey_f32"]).max()
    assert d > 1.0


@pytest.mark.parametrize("win", ["hamming", "povey"])
def test_config1_10s(frontend_golden, win):
    g = frontend_golden
    import torch

    gen = torch.Generator().manual_seed(1234)
    w = (0.5 * (2 * torch.rand(160000, generator=gen) - 1)).numpy()
    feats, lens = of.wav_frontend([w], win)
    assert feats.shape == (1, 167, 560) and lens[0] == 167
    assert np.abs(feats[0, ::8] - g[f"cfg1_{win}_rows"]).max() < 1e-3
    np.testing.assert_allclose(feats[0].astype(np.float64).sum(0), g[f"cfg1_{win}_colsum"], atol=2e-2)


def test_short_waveform_raises():
    with pytest.raises(ValueError):
        of.fbank(np.zeros(399, np.float32))
