"""The drop-in seam itself: ``install()`` into the class registry the reference resolves names in
(/root/reference/model.py:57-59,112-118) and ``patch(model, kwargs)`` on a built model (/root/reference/infer.py:377-392,
455-458).  funasr is not on this box; the stand-in model carries ``oracle.encoder`` modules, which use funasr's attribute
and parameter names (``encoders0[0].self_attn.h``, ``.fsmn_block``, ``feed_forward.w_1``, ``blocks[0].self_attn.h``,
``linear1`` ...), and the frontend is vLLM's verbatim copy of funasr's ``WavFrontend`` where importable."""
import types

import numpy as np
import pytest
import torch


def _stand_in(n_blocks=2, tp_blocks=1, sanm_shift=0):
    from oracle import encoder as oe

    conf = dict(oe.NANO_ENCODER_CONF, num_blocks=n_blocks, tp_blocks=tp_blocks, sanm_shift=sanm_shift)
    m = types.SimpleNamespace()
    m.audio_encoder = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **conf), seed=31)
    m.audio_adaptor = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=32)
    return m


class _FunasrStyleFrontend:
    """Attribute surface of funasr's WavFrontend (VLLM/transformers_utils/processors/funasr.py:89-123)."""

    def __init__(self, cmvn=None, dither=1.0, window="hamming"):
        self.fs, self.window, self.n_mels, self.frame_length, self.frame_shift = 16000, window, 80, 25, 10
        self.filter_length_min = self.filter_length_max = -1
        self.lfr_m, self.lfr_n, self.cmvn_file, self.dither = 7, 6, None, dither
        self.snip_edges, self.upsacle_samples, self.cmvn = True, True, cmvn


def test_conf_is_recovered_from_a_funasr_style_module_tree():
    from funasr_b200 import registry

    m = _stand_in(3, 2, sanm_shift=2)
    c = registry._conf_of_encoder(m.audio_encoder)
    assert c == dict(input_size=560, output_size=512, attention_heads=4, linear_units=2048, num_blocks=3, tp_blocks=2,
                     kernel_size=11, sanm_shfit=2)
    a = registry._conf_of_adaptor(m.audio_adaptor)
    assert a == dict(downsample_rate=1, encoder_dim=512, llm_dim=1024, ffn_dim=2048, n_layer=2, attention_heads=8)

    class _Pad:  # funasr keeps the left pad in a ConstantPad1d (VLLM/model_executor/models/funasr.py:175-179)
        padding = (8, 2)

    attn = types.SimpleNamespace(pad_fn=_Pad())
    assert registry._sanm_shift_of(attn, 11) == 3
    assert registry._sanm_shift_of(types.SimpleNamespace(), 11) == 0


def test_install_resolves_the_names_the_reference_looks_up():
    from funasr_b200 import registry
    import funasr_b200 as fb

    t = registry.install(override=False)
    assert t.encoder_classes.get("SenseVoiceEncoderSmallB200") is fb.SenseVoiceEncoderSmall
    assert t.adaptor_classes.get("TransformerB200") is fb.Transformer
    assert t.frontend_classes.get("WavFrontendB200") is fb.WavFrontend
    t = registry.install(override=True)
    # model.py:57: tables.encoder_classes.get(audio_encoder)(input_size=..., **audio_encoder_conf); :59 .output_size()
    enc = t.encoder_classes.get("SenseVoiceEncoderSmall")(input_size=560, output_size=512, attention_heads=4, linear_units=2048,
                                                          num_blocks=2, tp_blocks=1, kernel_size=11, sanm_shfit=0,
                                                          dropout_rate=0.1, selfattention_layer_type="sanm")
    assert enc.output_size() == 512
    # model.py:112-118: adaptor_conf gets encoder_dim / llm_dim injected
    ad = t.adaptor_classes.get("Transformer")(downsample_rate=1, encoder_dim=512, llm_dim=1024, ffn_dim=2048, n_layer=2)
    assert ad.linear2.weight.shape == (1024, 2048)
    # an unchanged frontend_conf has no dither key -> funasr's default 1.0: must build (and say that dither is ignored)
    with pytest.warns(UserWarning, match="dither"):
        import funasr_b200.frontend as ff

        ff._dither_warned = False
        fe = t.frontend_classes.get("WavFrontend")(cmvn_file=None, fs=16000, window="hamming", n_mels=80, frame_length=25,
                                                   frame_shift=10, lfr_m=7, lfr_n=6)
    assert fe.output_size() == 560 and fe.dither == 1.0


def test_load_cmvn_matches_the_reference_parser(frontend_golden, tmp_path):
    """frontend.load_cmvn vs the committed am.mvn text and the table vLLM's own load_cmvn returned for it."""
    from funasr_b200.frontend import load_cmvn
    from oracle import frontend as of

    text = str(frontend_golden["cmvn_text"])
    p = tmp_path / "am.mvn"
    p.write_text(text, encoding="utf-8")
    got = load_cmvn(str(p))
    want = of.parse_cmvn_text(text)
    assert got.dtype == torch.float32 and tuple(got.shape) == (2, 560)
    assert np.array_equal(got.numpy(), want)
    try:
        from vllm.transformers_utils.processors.funasr import load_cmvn as ref_load_cmvn
    except Exception:
        return
    assert torch.equal(got, ref_load_cmvn(str(p)))  # the reference-side parser itself, where importable


@pytest.mark.gpu
def test_patch_swaps_the_three_seams_and_keeps_the_weights(frontend_golden, tmp_path):
    import funasr_b200 as fb
    from funasr_b200 import registry
    from oracle import frontend as of

    m = _stand_in(2, 1)
    o_enc, o_ad = m.audio_encoder, m.audio_adaptor
    text = str(frontend_golden["cmvn_text"])
    cmvn_path = tmp_path / "am.mvn"
    cmvn_path.write_text(text, encoding="utf-8")
    try:  # vLLM's verbatim copy of funasr's WavFrontend = what AutoModel.build_model puts in kwargs["frontend"]
        from vllm.transformers_utils.processors.funasr import WavFrontend as RefFrontend

        old_fe = RefFrontend(cmvn_file=str(cmvn_path), fs=16000, window="hamming", n_mels=80, frame_length=25, frame_shift=10,
                             lfr_m=7, lfr_n=6, dither=0.0)
    except Exception:
        old_fe = _FunasrStyleFrontend(cmvn=torch.from_numpy(of.parse_cmvn_text(text)), dither=0.0)
    kwargs = {"frontend": old_fe, "tokenizer": object(), "device": "cuda"}
    tower = registry.patch(m, kwargs, device="cuda")
    assert isinstance(m.audio_encoder, fb.SenseVoiceEncoderSmall) and isinstance(m.audio_adaptor, fb.Transformer)
    assert isinstance(kwargs["frontend"], fb.WavFrontend) and m.audio_tower_b200 is tower
    for k, v in o_enc.state_dict().items():  # strict load happened: same keys, same values
        assert torch.equal(m.audio_encoder.state_dict()[k].cpu(), v), k
    for k, v in o_ad.state_dict().items():
        assert torch.equal(m.audio_adaptor.state_dict()[k].cpu(), v), k
    # the reference's batched loop through the patched seams (infer.py:446-466)
    g = torch.Generator().manual_seed(5)
    wavs = [0.5 * (2 * torch.rand(n, generator=g) - 1) for n in (32000, 8000, 16007)]
    speech, lens = fb.extract_fbank(wavs, frontend=kwargs["frontend"], is_final=True)
    y, ol = m.audio_encoder(speech, lens)
    z, zl = m.audio_adaptor(y, ol)
    tables = {"window": frontend_golden["window_hamming"], "mel": frontend_golden["mel_banks"]}
    want_f, want_l = of.wav_frontend([w.numpy() for w in wavs], "hamming", of.parse_cmvn_text(text), tables=tables)
    assert lens.tolist() == want_l.tolist() == zl.tolist()
    assert float((speech.cpu() - torch.from_numpy(want_f)).abs().max()) < 1e-3
    want_y, _ = o_enc(torch.from_numpy(want_f), torch.from_numpy(want_l))
    want_z, _ = o_ad(want_y, torch.from_numpy(want_l))
    rel = lambda a, b: float((a.double().cpu() - b.double()).pow(2).mean().sqrt() / b.double().pow(2).mean().sqrt())
    for b, t in enumerate(want_l.tolist()):
        assert rel(z[b, :t], want_z[b, :t]) < 2e-2
    # patching twice is a no-op on already swapped objects
    again = registry.patch(m, kwargs, device="cuda")
    assert isinstance(again, fb.AudioTower) and isinstance(kwargs["frontend"], fb.WavFrontend)


@pytest.mark.gpu
def test_patch_keeps_a_non_zero_sanm_shift():
    """An encoder built with sanm_shfit != 0 must be rebuilt with the same FSMN left pad."""
    from funasr_b200 import registry

    m = _stand_in(2, 0, sanm_shift=3)
    o_enc = m.audio_encoder
    registry.patch(m, None, device="cuda")
    assert m.audio_encoder.sanm_shift == 3
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 40, 560, generator=g)
    il = torch.tensor([40, 23])
    x[1, 23:] = 0
    want, _ = o_enc(x, il)
    got, _ = m.audio_encoder(x.cuda(), il)
    rel = lambda a, b: float((a.double().cpu() - b.double()).pow(2).mean().sqrt() / b.double().pow(2).mean().sqrt())
    assert rel(got[0], want[0]) < 2e-2 and rel(got[1, :23], want[1, :23]) < 2e-2


@pytest.mark.gpu
def test_wavfrontend_forward_padded_batch(frontend_golden):
    """WavFrontend.forward(input [B, Nmax], input_lengths) - the padded form funasr's extract_fbank hands the frontend
    (VLLM:311-313) - against the committed outputs of the real WavFrontend + torchaudio."""
    import funasr_b200 as fb

    g = frontend_golden
    wavs = [torch.from_numpy(g[f"wav_{i}"]) for i in range(8)]
    lens = [int(w.numel()) for w in wavs]
    padded = torch.nn.utils.rnn.pad_sequence(wavs, batch_first=True, padding_value=0.0)  # VLLM:311
    for window in ("hamming", "povey"):
        fe = fb.WavFrontend(fs=16000, window=window, n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                            device="cuda")
        for lens_arg in (lens, torch.tensor(lens)):
            feats, flens = fe(padded, lens_arg)
            assert flens.tolist() == g[f"lens_{window}_nocmvn"].tolist()
            want = torch.from_numpy(g[f"feats_{window}_nocmvn"])
            assert tuple(feats.shape) == tuple(want.shape)
            assert float((feats.cpu() - want).abs().max()) < 1e-3
            for b, t in enumerate(flens.tolist()):
                assert torch.all(feats[b, t:] == 0)
    # a 1-D input is one utterance (VLLM:299-301)
    f1, l1 = fe(wavs[3], [lens[3]])
    assert tuple(f1.shape) == (1, 17, 560) and l1.tolist() == [17]
