"""The vLLM-native multimodal route: ``FunASRForConditionalGeneration.embed_multimodal``
(VLLM/model_executor/models/funasr.py:967-976 -> 694-713) replaced by the CUDA tower.

Expected values (``tests/golden/vllm_route_golden.npz``, generator ``make_golden_vllm_route.py``) are outputs of the stock
vLLM method on vLLM's own audio modules; the parameter names / shapes in the fixture are those of the real classes at full
Fun-ASR-Nano depth.  The CPU tests pin the name mapping and the host logic; the GPU test is the parity test."""
import os
import types

import numpy as np
import pytest
import torch
import torch.nn as nn

from conftest import GOLDEN


@pytest.fixture(scope="module")
def route_golden():
    return np.load(os.path.join(GOLDEN, "vllm_route_golden.npz"))


class _MetaModule:
    """Only ``state_dict()``: the names and shapes of a real vLLM module, tensors on the meta device."""

    def __init__(self, names, shapes):
        self._sd = {str(n): torch.empty([int(d) for d in str(s).split(",") if d], device="meta") for n, s in zip(names, shapes)}

    def state_dict(self):
        return dict(self._sd)


def test_name_mapping_covers_every_parameter_of_the_real_vllm_modules(route_golden):
    """vLLM names -> funasr names must be EXACTLY the state dict of the CUDA classes at full depth (strict load)."""
    from funasr_b200 import SenseVoiceEncoderSmall, Transformer
    from funasr_b200.vllm_route import funasr_state_dict
    from oracle import encoder as oe

    g = route_golden
    conf = dict(oe.NANO_ENCODER_CONF)
    conf["sanm_shfit"] = conf.pop("sanm_shift")
    with torch.device("meta"):
        enc = SenseVoiceEncoderSmall(560, **conf)
        ad = Transformer(**oe.NANO_ADAPTOR_CONF)
    for tag, mod in (("enc", enc), ("ad", ad)):
        mapped = funasr_state_dict(_MetaModule(g[f"names_{tag}"], g[f"shapes_{tag}"]))
        want = {k: tuple(v.shape) for k, v in mod.state_dict().items()}
        got = {k: tuple(v.shape) for k, v in mapped.items()}
        assert got == want, (sorted(set(got) ^ set(want))[:6], tag)
    assert any(".self_attn.qkv." in str(n) for n in g["names_ad"])  # the stacked projection is really there
    assert any(".out_proj." in str(n) for n in g["names_enc"])


class _StubPlan:
    device = torch.device("cpu")


class _StubTower:
    """Host-logic stand-in: 'embedding' of a feature row = its first two features and its row index within the item."""

    class _Enc:
        input_size = 4

    encoder = _Enc()

    def plan(self):
        return _StubPlan()

    def _emb(self, rows):
        return torch.stack([rows[:, 0], rows[:, 1], torch.arange(rows.shape[0], dtype=torch.float32)], dim=1)

    def forward_export(self, speech, lens, out_dtype=torch.float32):
        out = torch.zeros(speech.shape[0], speech.shape[1], 3, dtype=out_dtype)
        for b, n in enumerate(lens):
            out[b, : int(n)] = self._emb(speech[b, : int(n)].float()).to(out_dtype)
        return out, torch.as_tensor(lens, dtype=torch.int32)

    def forward_packed(self, feats, lens, out_dtype=torch.float32, **kw):
        cu = np.zeros(len(lens) + 1, np.int64)
        cu[1:] = np.cumsum(lens)
        out = torch.cat([self._emb(feats[cu[b]: cu[b + 1]].float()) for b in range(len(lens))]).to(out_dtype)
        return out, cu


def test_route_host_logic_padded_ragged_list_and_empty():
    from funasr_b200.vllm_route import MultimodalRoute

    g = torch.Generator().manual_seed(0)
    x = torch.randn(3, 7, 4, generator=g)
    lens, take = torch.tensor([7, 3, 5]), torch.tensor([4, 1, 3])
    padded = MultimodalRoute(_StubTower())(input_features=x, speech_lengths=lens, fake_token_lengths=take)
    assert tuple(padded.shape) == (3, 7, 3)  # the stock return: one padded tensor (VLLM/...:967-976)
    assert torch.equal(padded[1, 3:], torch.zeros(4, 3))
    rag = MultimodalRoute(_StubTower(), ragged=True)(input_features=x, speech_lengths=lens, fake_token_lengths=take)
    assert [tuple(r.shape) for r in rag] == [(4, 3), (1, 3), (3, 3)]
    for b in range(3):
        assert torch.equal(rag[b], padded[b, : int(take[b])])
    # vLLM hands batched fields over as lists when the items do not stack: per-item [1, T, d] / [T, d] tensors and 0-d lengths
    items = [x[0:1, :7], x[1, :3], x[2:3]]
    lst = MultimodalRoute(_StubTower())(input_features=items, speech_lengths=[torch.tensor(7), torch.tensor([3]), torch.tensor(5)],
                                        fake_token_lengths=[torch.tensor(4), torch.tensor(1), torch.tensor(3)])
    for b in range(3):
        assert torch.equal(lst[b], rag[b])
    assert MultimodalRoute(_StubTower())(input_features=None) == []
    from funasr_b200._lib import FunASRB200Error

    with pytest.raises(FunASRB200Error):
        MultimodalRoute(_StubTower())(input_features=x[0], speech_lengths=lens[:1])


def test_install_wraps_load_weights_of_the_engine_class(monkeypatch):
    vf = pytest.importorskip("vllm.model_executor.models.funasr")
    from funasr_b200 import vllm_route

    cls = vf.FunASRForConditionalGeneration
    calls = []
    monkeypatch.setattr(cls, "load_weights", lambda self, weights: {"stock"})
    monkeypatch.setattr(vllm_route, "patch_vllm_model", lambda m, **kw: calls.append((m, kw)))
    monkeypatch.setattr(vllm_route, "_INSTALLED", False)
    assert vllm_route.install_vllm_route(max_tokens=4096) is cls
    me = object()
    assert cls.load_weights(me, []) == {"stock"}
    assert calls == [(me, dict(max_tokens=4096, ragged=False))]
    vllm_route.install_vllm_route()  # idempotent: no second wrapper
    calls.clear()
    cls.load_weights(me, [])
    assert len(calls) == 1


# ----------------------------------------------------------------------------- GPU parity
def _vllm_style_modules():
    """Modules with vLLM's parameter names and attribute surface, carrying the golden's seeded weights.  Built from the
    oracle classes (funasr names) by the inverse renaming, so the route's own mapping is what brings them back."""
    from oracle import encoder as oe

    conf = dict(oe.NANO_ENCODER_CONF, num_blocks=2, tp_blocks=1)
    enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **conf), seed=4321)
    ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)
    for layer in list(enc.encoders0) + list(enc.encoders) + list(enc.tp_encoders):
        a = layer.self_attn
        a.out_proj = a.linear_out
        del a.linear_out
        a.pad_fn = nn.ConstantPad1d(a.pad, 0.0)  # VLLM/...:175-179
        del a.pad
    for blk in ad.blocks:
        a = blk.self_attn
        d = a.linear_q.weight.shape[0]
        a.qkv = nn.Linear(d, 3 * d)
        with torch.no_grad():
            a.qkv.weight.copy_(torch.cat([a.linear_q.weight, a.linear_k.weight, a.linear_v.weight]))
            a.qkv.bias.copy_(torch.cat([a.linear_q.bias, a.linear_k.bias, a.linear_v.bias]))
        a.out_proj, a.num_heads = a.linear_out, a.h
        del a.linear_q, a.linear_k, a.linear_v, a.linear_out, a.h
    return enc, ad


def test_stand_in_carries_the_real_vllm_names(route_golden):
    enc, ad = _vllm_style_modules()
    assert set(enc.state_dict()) <= {str(n) for n in route_golden["names_enc"]}
    assert set(ad.state_dict()) == {str(n) for n in route_golden["names_ad"]}


@pytest.mark.gpu
def test_route_matches_stock_embed_multimodal(route_golden):
    from funasr_b200.vllm_route import patch_vllm_model
    from gpu_util import rel_rms

    g = route_golden
    enc, ad = _vllm_style_modules()
    model = nn.Module()
    model.model = nn.Module()
    model.model.encoder = nn.Module()
    model.model.encoder.audio_encoder, model.model.encoder.audio_adaptor = enc.cuda(), ad.cuda()
    model.dtype = torch.float32
    tower = patch_vllm_model(model, max_tokens=256)
    assert model.audio_tower_b200 is tower

    x, lens, fake = torch.from_numpy(g["x"]), torch.from_numpy(g["lens"]), torch.from_numpy(g["fake_token_lengths"])
    B = x.shape[0]
    worst = 0.0
    singles = []
    for b in range(B):  # the form the fixture pins: one item per call, features padded to a fixed length
        y = model.embed_multimodal(input_features=x[b:b + 1].cuda(), speech_lengths=lens[b:b + 1], fake_token_lengths=fake[b:b + 1])
        assert tuple(y.shape) == (1, x.shape[1], 1024) and y.dtype == torch.float32
        t = int(lens[b])
        assert torch.count_nonzero(y[0, t:]) == 0
        worst = max(worst, rel_rms(y[0, :t], torch.from_numpy(g[f"embed_{b}"])))
        singles.append(y[0, :t].clone())
    print(f"vLLM-native route vs stock embed_multimodal: worst rel-rms {worst:.2e}")
    assert worst <= 2e-2  # bf16 operands against vLLM's fp32 modules (same bar as the encoder tests)

    # one batched call = the single-item calls, bit for bit (every item is computed alone)
    yb = model.embed_multimodal(input_features=x.cuda(), speech_lengths=lens.cuda(), fake_token_lengths=fake)
    for b in range(B):
        assert torch.equal(yb[b, : int(lens[b])], singles[b])
    # ragged result: the rows the model runner keeps per item (fake_token_len of them)
    model.dtype = torch.bfloat16
    patch_vllm_model(model, max_tokens=256, ragged=True)
    rag = model.embed_multimodal(input_features=x.cuda(), speech_lengths=lens, fake_token_lengths=fake)
    lst = model.embed_multimodal(input_features=[x[b, : int(lens[b])].cuda() for b in range(B)], speech_lengths=list(lens),
                                 fake_token_lengths=list(fake))
    for b in range(B):
        assert tuple(rag[b].shape) == (int(fake[b]), 1024) and rag[b].dtype == torch.bfloat16
        assert torch.equal(rag[b], lst[b])
        assert rel_rms(rag[b].float(), singles[b][: int(fake[b])]) <= 1e-2  # bf16 output rounding only
