"""Pins oracle/encoder.py to vLLM's CPU-built SenseVoiceEncoderSmall / adaptor outputs (CPU)."""
import ast

import numpy as np
import torch

from oracle import encoder as oe


def _sd(g, prefix):
    return {k[len(prefix):]: torch.from_numpy(g[k]) for k in g.files if k.startswith(prefix)}


def test_small_encoder_vs_vllm(encoder_golden):
    g = encoder_golden
    cfg = ast.literal_eval(str(g["A_cfg"]))
    enc = oe.SenseVoiceEncoderSmall(**cfg).eval()
    enc.load_state_dict(_sd(g, "A_w/"), strict=True)
    x, lens = torch.from_numpy(g["A_x"]), torch.from_numpy(g["A_lens"])
    y, ol = enc(x, lens)  # per-utterance oracle definition
    assert np.array_equal(ol.numpy(), g["A_olens"])
    for b in range(3):
        t = int(lens[b])
        np.testing.assert_allclose(y[b, :t].numpy(), g[f"A_y_single_{b}"], atol=2e-5, rtol=1e-5)
        # padded+masked batch == per-utterance on valid rows
        np.testing.assert_allclose(y[b, :t].numpy(), g["A_y_pad"][b, :t], atol=5e-5, rtol=1e-5)
    yp, olp = enc.forward_padded(x, lens)
    np.testing.assert_allclose(yp.numpy()[0], g["A_y_pad"][0], atol=5e-5, rtol=1e-5)
    for b in range(3):
        t = int(lens[b])
        np.testing.assert_allclose(yp.numpy()[b, :t], g["A_y_pad"][b, :t], atol=5e-5, rtol=1e-5)


def test_small_adaptor_vs_vllm(encoder_golden):
    g = encoder_golden
    cfg = ast.literal_eval(str(g["A_ad_cfg"]))
    ad = oe.Transformer(**cfg).eval()
    ad.load_state_dict(_sd(g, "A_adw/"), strict=True)
    for b in range(3):
        y = torch.from_numpy(g[f"A_y_single_{b}"])[None]
        z, ol = ad(y, torch.tensor([y.shape[1]]))
        assert int(ol[0]) == y.shape[1]
        np.testing.assert_allclose(z[0].numpy(), g[f"A_z_single_{b}"], atol=5e-5, rtol=1e-5)


def test_nano_width_vs_vllm(encoder_golden):
    g = encoder_golden
    conf = dict(oe.NANO_ENCODER_CONF, num_blocks=2, tp_blocks=1)
    enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **conf), seed=4321)
    ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)
    x, lens = torch.from_numpy(g["B_x"]), torch.from_numpy(g["B_lens"])
    y, _ = enc(x, lens)
    for b in range(2):
        t = int(lens[b])
        np.testing.assert_allclose(y[b, :t].numpy(), g["B_y_pad"][b, :t], atol=1e-4, rtol=1e-4)
        z, _ = ad(torch.from_numpy(g["B_y_pad"][b : b + 1, :t]), lens[b : b + 1])
        np.testing.assert_allclose(z[0].numpy(), g[f"B_z_single_{b}"], atol=2e-4, rtol=1e-4)


def test_layer0_has_no_attention_residual():
    # in_size != size -> x = attn(LN(x)), no "+ x"
    torch.manual_seed(0)
    l0 = oe.SanmLayer(48, 32, 4, 64, 11).eval()
    x = torch.randn(1, 7, 48)
    a = l0.self_attn(l0.norm1(x))
    want = a + l0.feed_forward(l0.norm2(a))
    torch.testing.assert_close(l0(x), want)


def test_pe_closed_form():
    # SURVEY.md §8a7
    pe = oe.sinusoidal_pe(5, 560)
    k = torch.arange(280, dtype=torch.float32)
    inv = torch.exp(k * (-(torch.log(torch.tensor([10000.0])) / 279)))
    for p in range(1, 6):
        torch.testing.assert_close(pe[p - 1, :280], torch.sin(p * inv))
        torch.testing.assert_close(pe[p - 1, 280:], torch.cos(p * inv))


def test_param_count_matches_survey():
    enc = oe.SenseVoiceEncoderSmall(560, **oe.NANO_ENCODER_CONF)
    ad = oe.Transformer(**oe.NANO_ADAPTOR_CONF)
    n = sum(p.numel() for p in enc.parameters()) + sum(p.numel() for p in ad.parameters())
    assert abs(n / 1e6 - 233.7) < 0.5
