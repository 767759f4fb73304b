"""The vLLM-native multimodal route: the CUDA audio path INSIDE the engine.

With ``enable_prompt_embeds`` the reference computes the embeddings in its own process and ships them to the engine
through ``.cpu()`` + msgpack (``/root/reference/infer.py:462-475``, VLLM/renderers/base.py:755-758).  vLLM's own FunASR
model takes the other road: the request carries LFR/CMVN features and the engine computes the embeddings in
``FunASRForConditionalGeneration.embed_multimodal`` (VLLM/model_executor/models/funasr.py:967-976) through
``FunASRModel.get_encoder_outputs`` (…:694-713) = ``audio_encoder`` -> ``audio_adaptor``.  This module replaces THAT
call with the packed-row CUDA tower, so the embeddings are born in the engine's HBM and never cross the host:

  * ``funasr_state_dict(module)``  vLLM parameter names -> funasr checkpoint names (the inverse of the engine's
    ``hf_to_vllm_mapper`` / ``stacked_params_mapping``, …:641-646, 845-856): ``out_proj`` -> ``linear_out``, the adaptor's
    stacked ``qkv`` -> ``linear_q`` / ``linear_k`` / ``linear_v``.
  * ``patch_vllm_model(model)``   builds the CUDA ``SenseVoiceEncoderSmall`` / ``Transformer`` from
    ``model.model.encoder.audio_encoder`` / ``.audio_adaptor`` (weights copied, strict), and binds a new
    ``embed_multimodal`` on the instance.  Same keyword inputs (``input_features`` ``[B, T, 560]``, ``speech_lengths``
    ``[B]``, ``fake_token_lengths`` ``[B]``), same result: ``[B, T, llm_dim]`` in the model dtype.  With
    ``ragged=True`` the result is a tuple of per-item ``[fake_token_len_b, llm_dim]`` row views instead (what the model
    runner slices out of the padded tensor anyway, and what ``/root/reference/model.py:513-541`` keeps).
  * ``install_vllm_route()``      class-level seam for engine processes: wraps ``load_weights`` of
    ``FunASRForConditionalGeneration`` so every instance is patched right after its checkpoint is in (call it from a
    ``vllm.general_plugins`` entry point).

One stated difference: vLLM's adaptor blocks attend over the zero padding of a batch (``EncoderLayer.forward`` passes
no mask, …:447-455), so a padded multi-item call of the stock method is not equal to its own single-item calls.  The CUDA
tower computes every item alone (funasr's masked semantics, what the reference process computes), so it matches the
stock method item by item on single-item calls - the form the parity test pins.

Tensor parallelism: the audio tower is replicated, not sharded; a TP > 1 engine shards these weights and is refused.
"""
from __future__ import annotations

import types

import numpy as np
import torch

from . import _lib
from .encoder import AudioTower, SenseVoiceEncoderSmall, Transformer
from .registry import _conf_of_encoder


def funasr_state_dict(module) -> dict:
    """State dict of a vLLM ``SenseVoiceEncoderSmall`` / adaptor ``Transformer`` under funasr's parameter names."""
    out = {}
    for name, t in module.state_dict().items():
        name = name.replace(".out_proj.", ".linear_out.")
        if ".self_attn.qkv." in name:  # QKVParallelLinear: rows are [q | k | v] (…:479-486, 641-646)
            if t.shape[0] % 3:
                raise _lib.FunASRB200Error(f"{name}: stacked qkv with {t.shape[0]} rows")
            for part, rows in zip(("linear_q", "linear_k", "linear_v"), t.chunk(3, dim=0)):
                out[name.replace(".qkv.", f".{part}.")] = rows
        else:
            out[name] = t
    return out


def _adaptor_conf(ad) -> dict:
    blocks = getattr(ad, "blocks", None)
    n_layer = len(blocks) if blocks is not None else 0
    sa = blocks[0].self_attn if n_layer else None
    heads = (getattr(sa, "num_heads", None) or getattr(sa, "h", None) or 8) if sa is not None else 8
    k = int(ad.k)
    return dict(downsample_rate=k, encoder_dim=ad.linear1.weight.shape[1] // k, llm_dim=ad.linear2.weight.shape[0],
                ffn_dim=ad.linear1.weight.shape[0], n_layer=n_layer, attention_heads=int(heads))


def _tp_world_size() -> int:
    try:
        from vllm.distributed import get_tensor_model_parallel_world_size

        return int(get_tensor_model_parallel_world_size())
    except Exception:
        return 1  # no vLLM / no TP group: nothing is sharded


@torch.no_grad()
def build_tower(audio_encoder, audio_adaptor, device=None, max_tokens: int = 8192) -> AudioTower:
    """CUDA tower carrying the weights of vLLM's two audio modules."""
    if _tp_world_size() != 1:
        raise _lib.FunASRB200Error("the audio tower is replicated per GPU; a tensor-parallel engine shards its weights")
    device = torch.device(device) if device is not None else next(audio_encoder.parameters()).device
    if device.type != "cuda":
        device = torch.device("cuda", torch.cuda.current_device())
    enc = SenseVoiceEncoderSmall(**_conf_of_encoder(audio_encoder))
    ad = Transformer(**_adaptor_conf(audio_adaptor))
    enc.load_state_dict({k: v.float() for k, v in funasr_state_dict(audio_encoder).items()}, strict=True)
    ad.load_state_dict({k: v.float() for k, v in funasr_state_dict(audio_adaptor).items()}, strict=True)
    enc.to(device).eval()
    ad.to(device).eval()
    return AudioTower(enc, ad, max_tokens=max_tokens)


def _as_lens(v, n_items: int | None = None) -> np.ndarray | None:
    if v is None:
        return None
    if isinstance(v, (list, tuple)):
        v = [int(torch.as_tensor(x).reshape(-1)[0]) for x in v]
        return np.asarray(v, dtype=np.int64)
    return torch.as_tensor(v).detach().reshape(-1).to("cpu").numpy().astype(np.int64)


class MultimodalRoute:
    """``embed_multimodal`` of one patched model (callable with the stock keyword arguments)."""

    def __init__(self, tower: AudioTower, out_dtype=torch.float32, ragged: bool = False):
        self.tower, self.out_dtype, self.ragged = tower, out_dtype, ragged

    @torch.no_grad()
    def __call__(self, **kwargs):
        feats = kwargs.get("input_features")
        lens = _as_lens(kwargs.get("speech_lengths"))
        take = _as_lens(kwargs.get("fake_token_lengths"))
        if feats is None:
            return []  # no audio item in this call (the stock schema allows None, …:607-618)
        dev = self.tower.plan().device
        if isinstance(feats, (list, tuple)):  # ragged items: pack the rows, no padded tensor on either side
            items = [f.reshape(-1, f.shape[-1]) for f in feats]
            if lens is None:
                lens = np.asarray([f.shape[0] for f in items], dtype=np.int64)
            packed = torch.cat([f[: int(n)] for f, n in zip(items, lens)], dim=0).to(dev, torch.float32).contiguous()
            out, cu = self.tower.forward_packed(packed, lens, out_dtype=self.out_dtype)
            rows = [out[int(cu[b]): int(cu[b + 1])] for b in range(len(lens))]
            if take is not None:
                rows = [r[: int(t)] for r, t in zip(rows, take)]
            return tuple(rows)
        if feats.dim() != 3:
            raise _lib.FunASRB200Error(f"input_features must be [B, T, {self.tower.encoder.input_size}], got {tuple(feats.shape)}")
        if lens is None:
            lens = np.full(feats.shape[0], feats.shape[1], dtype=np.int64)
        out, _ = self.tower.forward_export(feats.to(dev), lens, out_dtype=self.out_dtype)
        if self.ragged:
            keep = lens if take is None else np.minimum(take, lens)
            return tuple(out[b, : int(keep[b])] for b in range(out.shape[0]))
        return out


def patch_vllm_model(model, max_tokens: int = 8192, ragged: bool = False, device=None) -> AudioTower:
    """Swap the audio path of a built vLLM ``FunASRForConditionalGeneration`` (…:842-1013) for the CUDA tower, in place."""
    enc_holder = model.model.encoder
    tower = build_tower(enc_holder.audio_encoder, enc_holder.audio_adaptor, device=device, max_tokens=max_tokens)
    dtype = getattr(model, "dtype", None)
    if not isinstance(dtype, torch.dtype) or dtype not in (torch.float32, torch.bfloat16):
        dtype = torch.float32
    route = MultimodalRoute(tower, out_dtype=dtype, ragged=ragged)
    model.audio_tower_b200 = tower
    model.embed_multimodal = types.MethodType(lambda self, **kw: route(**kw), model)
    return tower


_INSTALLED = False


def install_vllm_route(max_tokens: int = 8192, ragged: bool = False):
    """Patch every ``FunASRForConditionalGeneration`` the engine builds, right after its weights are loaded."""
    global _INSTALLED
    from vllm.model_executor.models.funasr import FunASRForConditionalGeneration as cls

    if _INSTALLED:
        return cls
    stock_load = cls.load_weights

    def load_weights(self, weights):
        loaded = stock_load(self, weights)
        patch_vllm_model(self, max_tokens=max_tokens, ragged=ragged)
        return loaded

    cls.load_weights = load_weights
    _INSTALLED = True
    return cls
