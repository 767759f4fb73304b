"""The drop-in seams of the reference.

(1) class registry — the reference resolves the encoder / adaptor by name in ``funasr.register.tables``
    (``/root/reference/model.py:14,57-59,112-118``).  ``install()`` registers the CUDA classes there, under their own
    names (``SenseVoiceEncoderSmallB200``, ``TransformerB200``, ``WavFrontendB200``) and, with ``override=True``, under
    funasr's names so an unchanged ``config.yaml`` resolves to them.  When funasr is not importable (this box) the same
    calls work on a local ``tables`` object with funasr's interface, so host logic and tests do not need funasr.
(2) object attributes — ``patch(model, kwargs)`` swaps ``model.audio_encoder``, ``model.audio_adaptor`` and
    ``kwargs["frontend"]`` of an already built ``FunASRNano`` (``infer.py:377,392,455,458``) and copies the weights.
"""
from __future__ import annotations

import torch

from .encoder import AudioTower, SenseVoiceEncoderSmall, Transformer
from .frontend import WavFrontend


class _ClassTable(dict):
    pass


class _Tables:
    """Minimal stand-in for ``funasr.register.tables``: ``<kind>_classes`` dicts + ``register(kind, name)`` decorator."""

    def __init__(self):
        for kind in ("model_classes", "frontend_classes", "encoder_classes", "adaptor_classes", "tokenizer_classes"):
            setattr(self, kind, _ClassTable())

    def register(self, kind: str, name: str | None = None):
        def deco(cls):
            getattr(self, kind)[name or cls.__name__] = cls
            return cls

        return deco


def _funasr_tables():
    try:
        from funasr.register import tables as t  # type: ignore

        return t
    except Exception:
        return None


tables = _funasr_tables() or _Tables()


def install(override: bool = False):
    """Register the CUDA classes into the registry the reference looks classes up in.  Returns that registry."""
    reg = [("encoder_classes", "SenseVoiceEncoderSmallB200", SenseVoiceEncoderSmall),
           ("adaptor_classes", "TransformerB200", Transformer),
           ("frontend_classes", "WavFrontendB200", WavFrontend)]
    if override:
        reg += [("encoder_classes", "SenseVoiceEncoderSmall", SenseVoiceEncoderSmall),
                ("adaptor_classes", "Transformer", Transformer),
                ("frontend_classes", "WavFrontend", WavFrontend)]
    for kind, name, cls in reg:
        getattr(tables, kind)[name] = cls
    return tables


def _sanm_shift_of(attn, kernel_size: int) -> int:
    """funasr keeps the FSMN left pad in ``pad_fn = ConstantPad1d((left, right), 0)`` with ``left = (k-1)//2 + sanm_shfit``
    (VLLM/model_executor/models/funasr.py:175-179); recover the shift from it (the oracle classes keep ``pad``)."""
    pad = getattr(getattr(attn, "pad_fn", None), "padding", None) or getattr(attn, "pad", None)
    if pad is None:
        return 0
    return max(int(pad[0]) - (kernel_size - 1) // 2, 0)


def _conf_of_encoder(enc) -> dict:
    l0 = enc.encoders0[0]
    a = l0.self_attn
    k = a.fsmn_block.weight.shape[-1]
    return dict(input_size=l0.norm1.weight.numel(), output_size=l0.norm2.weight.numel(), attention_heads=a.h,
                linear_units=l0.feed_forward.w_1.weight.shape[0], num_blocks=1 + len(enc.encoders),
                tp_blocks=len(getattr(enc, "tp_encoders", [])), kernel_size=k, sanm_shfit=_sanm_shift_of(a, k))


def _conf_of_adaptor(ad) -> dict:
    blocks = getattr(ad, "blocks", None)
    n_layer = len(blocks) if blocks is not None else 0
    heads = blocks[0].self_attn.h if n_layer else 8
    return dict(downsample_rate=ad.k, encoder_dim=ad.linear1.weight.shape[1] // ad.k, llm_dim=ad.linear2.weight.shape[0],
                ffn_dim=ad.linear1.weight.shape[0], n_layer=n_layer, attention_heads=heads)


@torch.no_grad()
def patch(model, kwargs: dict | None = None, device=None, max_tokens: int = 4096):
    """Swap the audio path of a built ``FunASRNano`` for the CUDA one, in place.  The state dict of the old modules is
    loaded into the new ones (same key names), so outputs follow the same weights.  Returns the ``AudioTower``."""
    old_enc, old_ad = model.audio_encoder, model.audio_adaptor
    device = torch.device(device) if device is not None else next(old_enc.parameters()).device
    if device.type != "cuda":
        device = torch.device("cuda", torch.cuda.current_device())
    enc = old_enc if isinstance(old_enc, SenseVoiceEncoderSmall) else SenseVoiceEncoderSmall(**_conf_of_encoder(old_enc))
    ad = old_ad if isinstance(old_ad, Transformer) else Transformer(**_conf_of_adaptor(old_ad))
    if enc is not old_enc:
        enc.load_state_dict(old_enc.state_dict(), strict=True)
    if ad is not old_ad:
        ad.load_state_dict(old_ad.state_dict(), strict=True)
    enc.to(device).eval()
    ad.to(device).eval()
    model.audio_encoder, model.audio_adaptor = enc, ad
    tower = AudioTower(enc, ad, max_tokens=max_tokens)
    model.audio_tower_b200 = tower
    if kwargs is not None and kwargs.get("frontend") is not None and not isinstance(kwargs["frontend"], WavFrontend):
        f = kwargs["frontend"]
        if getattr(f, "dither", 0.0) != 0.0:
            from .frontend import _warn_dither

            _warn_dither(f.dither)  # the reference's default adds random dither; the CUDA frontend does not
        kwargs["frontend"] = WavFrontend(cmvn_file=getattr(f, "cmvn_file", None), fs=f.fs, window=f.window, n_mels=f.n_mels,
                                         frame_length=f.frame_length, frame_shift=f.frame_shift, lfr_m=f.lfr_m, lfr_n=f.lfr_n,
                                         dither=0.0, upsacle_samples=getattr(f, "upsacle_samples", True), device=device,
                                         cmvn=getattr(f, "cmvn", None))
    return tower
