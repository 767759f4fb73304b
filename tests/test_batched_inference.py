"""batched m.inference through the reference's own `audio_embedding` bypass (model.py:477-479).
The stand-in model below restates the parts of FunASRNano.inference_prepare that matter at this seam (model.py:463-541):
refuse batch_size > 1, run encoder + adaptor unless audio_embedding / audio_embedding_lens are given, take
`fake_token_len` rows (model.py:391-396) of the embedding.  The test checks that the batched call hands every utterance
the rows the per-utterance reference path would have computed itself."""
import numpy as np
import pytest
import torch


class _StandInNano:
    """Just the seam: what FunASRNano.inference does with the audio (model.py:463-541), no LLM."""

    def __init__(self, enc, ad, use_low_frame_rate=True):
        self.audio_encoder, self.audio_adaptor, self.use_low_frame_rate = enc, ad, use_low_frame_rate
        self.calls = []

    def inference(self, data_in, data_lengths=None, key=None, tokenizer=None, frontend=None, **kwargs):
        import funasr_b200 as fb

        if kwargs.get("batch_size", 1) > 1:
            raise NotImplementedError("batch decoding is not implemented")  # model.py:463-464
        speech, lens = fb.extract_fbank(data_in[0], frontend=frontend, is_final=True)  # model.py:372
        n = fb.fake_token_len(int(lens[0]), self.use_low_frame_rate)  # model.py:391-396
        used_bypass = "audio_embedding" in kwargs and "audio_embedding_lens" in kwargs  # model.py:477
        if used_bypass:
            out, olens = kwargs["audio_embedding"], kwargs["audio_embedding_lens"]
        else:
            y, ol = self.audio_encoder(speech, lens)
            out, olens = self.audio_adaptor(y, ol)
        self.calls.append(dict(bypass=used_bypass, key=key))
        return out[0, :n].float().cpu(), int(olens[0])  # model.py:518-519: encoder_out[speech_idx, :speech_token_len]


def test_argument_errors():
    from funasr_b200 import batched

    with pytest.raises(ValueError):
        batched.inference_batched(object(), [], frontend=None)
    with pytest.raises(ValueError):
        batched._as_waveform("a.wav", None)
    with pytest.raises(ValueError):
        batched._as_waveform(torch.zeros(2, 3), None)
    assert batched._as_waveform(np.zeros((1, 5), np.float32), None).shape == (5,)
    assert batched._as_waveform("x", lambda p: torch.ones(7)).shape == (7,)


def test_known_features_stand_in():
    """batched.KnownFeatures: the utterance it was built for gets its features back from funasr's call form
    (`frontend(data, data_len)`) and from this package's extract_fbank; anything else reaches the real frontend."""
    import funasr_b200 as fb
    from funasr_b200.batched import KnownFeatures

    class Real:
        fs, frame_shift, lfr_n = 16000, 10, 6
        calls = 0

        def __call__(self, x, lens=None, **kw):
            Real.calls += 1
            return "real", lens

    feats, lens = torch.arange(2 * 560, dtype=torch.float32).reshape(1, 2, 560), torch.tensor([2])
    kf = KnownFeatures(Real(), 1000, feats, lens)
    assert kf.fs == 16000 and kf.lfr_n == 6
    for x, l in ((torch.zeros(1, 1000), torch.tensor([1000])), (torch.zeros(1000), None), (np.zeros((1, 1000), np.float32), [1000])):
        f, n = kf(x, l)
        assert f is feats and n is lens
    assert kf(torch.zeros(1, 999), [999])[0] == "real" and kf(torch.zeros(2, 1000), [1000, 1000])[0] == "real" and Real.calls == 2
    f, n = fb.extract_fbank([torch.zeros(1000)], frontend=kf, is_final=True)
    assert torch.equal(f, feats) and n.dtype == torch.int32 and n.tolist() == [2]
    assert kf.known(1001) is None


@pytest.mark.gpu
def test_batched_inference_equals_per_utterance_reference_path():
    import funasr_b200 as fb
    from funasr_b200.batched import inference_batched
    from oracle import encoder as oe

    conf = dict(oe.NANO_ENCODER_CONF, num_blocks=2, tp_blocks=1)
    enc = fb.SenseVoiceEncoderSmall(560, **conf)
    ad = fb.Transformer(**oe.NANO_ADAPTOR_CONF)
    enc.load_state_dict(oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **conf), seed=11).state_dict())
    ad.load_state_dict(oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=12).state_dict())
    enc.cuda(), ad.cuda()
    fe = fb.WavFrontend(fs=16000, window="hamming", n_mels=80, frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0,
                        cmvn=None, device="cuda")
    model = _StandInNano(enc, ad)
    model.audio_tower_b200 = fb.AudioTower(enc, ad)
    g = torch.Generator().manual_seed(3)
    wavs = [0.5 * (2 * torch.rand(n, generator=g) - 1) for n in (16000, 52000, 8123, 160000)]
    with pytest.raises(NotImplementedError):
        model.inference(data_in=wavs, frontend=fe, batch_size=4)  # the reference's behaviour
    n_fbank = [0]
    real_forward_packed = fe.forward_packed

    def counting(*a, **k):
        n_fbank[0] += 1
        return real_forward_packed(*a, **k)

    fe.forward_packed = counting
    got = inference_batched(model, wavs, frontend=fe, batch_size=4, key=["a", "b", "c", "d"], embed_dtype=torch.float32)
    fe.forward_packed = real_forward_packed
    # one fbank launch for the whole batch: the model's own extract_fbank call per utterance (model.py:372) got the known features
    assert n_fbank[0] == 1, n_fbank
    assert all(c["bypass"] for c in model.calls) and [c["key"] for c in model.calls] == [["a"], ["b"], ["c"], ["d"]]
    for w, (rows, olen) in zip(wavs, got):
        want_rows, want_len = model.inference(data_in=[w], frontend=fe)  # per-utterance path: encoder + adaptor inside
        assert olen == want_len and rows.shape == want_rows.shape
        # batched path emits bf16 rows (the prompt_embeds dtype), the per-utterance modules fp32: bf16 rounding only
        rel = float((rows - want_rows).pow(2).mean().sqrt() / want_rows.pow(2).mean().sqrt())
        assert rel < 6e-3, rel
