"""Batched ``m.inference`` for the audio path.

The reference refuses ``batch_size > 1`` (``/root/reference/model.py:463-464``: ``NotImplementedError("batch decoding is not
implemented")``) and runs frontend, encoder and adaptor once per utterance.  ``FunASRNano.inference_prepare`` has a documented
bypass, though: when ``audio_embedding`` and ``audio_embedding_lens`` are passed, encoder and adaptor are skipped and the given
rows are spliced into the prompt (``model.py:477-479,513-541``; ``fake_token_len`` / ``use_low_frame_rate`` slicing stays the
model's own, ``model.py:391-396,518-519``).  ``inference_batched`` computes the embeddings of ALL utterances in one packed
pass (fbank -> SAN-M encoder -> adaptor, no padding) and then calls the unchanged ``model.inference`` once per utterance
with its slice, so the LLM side is exactly the reference's."""
from __future__ import annotations

from typing import Callable, Sequence

import numpy as np
import torch

from .encoder import AudioTower
from .frontend import WavFrontend, fake_token_len, pack_waveforms


def _as_waveform(x, loader: Callable | None):
    if isinstance(x, (str, bytes)) or hasattr(x, "__fspath__"):
        if loader is None:
            raise ValueError("data_in holds a path: pass loader=callable(path) -> 1-D float waveform at the frontend's rate "
                             "(the reference uses funasr's load_audio_text_image_video, model.py:362)")
        x = loader(x)
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(x)
    if not torch.is_tensor(x):
        raise TypeError(f"unsupported data_in item {type(x)}")
    x = x.to(torch.float32)
    if x.dim() == 2 and x.shape[0] == 1:
        x = x[0]
    if x.dim() != 1:
        raise ValueError(f"expected a 1-D waveform, got shape {tuple(x.shape)}")
    return x


class KnownFeatures:
    """Frontend stand-in for ONE utterance whose features the batched pass has already computed.  ``model.inference`` still
    calls ``extract_fbank(data_src, frontend=frontend)`` for every utterance (``/root/reference/model.py:372``: it needs the
    length for ``fake_token_len``); handed this object instead of the real frontend, that call gets the known ``[1, T, 560]``
    features back (funasr's ``extract_fbank`` calls ``frontend(data, data_len)``; this package's looks ``known`` up first)
    instead of running the fbank kernel a second time.  Anything else (another waveform, attribute reads such as ``fs``,
    ``frame_shift``, ``lfr_n``: ``model.py:379-384``) goes to the real frontend."""

    def __init__(self, real: WavFrontend, n_samples: int, feats: torch.Tensor, lens: torch.Tensor):
        self._real, self._n, self._feats, self._lens = real, int(n_samples), feats, lens

    def __getattr__(self, name):
        return getattr(self._real, name)

    def known(self, n_samples: int):
        return (self._feats, self._lens) if int(n_samples) == self._n else None

    def forward(self, input, input_lengths=None, **kwargs):
        x = input if torch.is_tensor(input) else torch.as_tensor(input)
        if input_lengths is None:
            lens = [int(x.shape[-1])]
        else:
            lens = [int(v) for v in (input_lengths.tolist() if torch.is_tensor(input_lengths) else input_lengths)]
        if (x.dim() == 1 or x.shape[0] == 1) and len(lens) == 1 and lens[0] == self._n:
            return self._feats, self._lens
        return self._real(input, input_lengths, **kwargs)

    __call__ = forward


@torch.no_grad()
def audio_embeddings(wavs: Sequence[torch.Tensor], frontend: WavFrontend, tower: AudioTower, out_dtype=torch.bfloat16,
                     return_features: bool = False):
    """One packed pass over a list of 1-D waveforms.  Returns (list of ``[T_b, llm_dim]`` GPU tensors - views of one packed
    buffer -, ``T`` int64 array of rows per utterance, ``fake`` int64 array: the rows ``model.py:391-396`` will take when
    ``use_low_frame_rate`` is set) and, with ``return_features``, the list of ``[T_b, 560]`` f32 feature views as a fourth item."""
    dev = tower.encoder._device()
    pcm, starts = pack_waveforms(list(wavs), dev)
    n = np.asarray([int(w.numel()) for w in wavs], np.int64)
    feats, cu = frontend.forward_packed(pcm, starts, n)
    out, _ = tower.forward_packed(feats, np.diff(cu))
    if out.dtype != out_dtype:
        out = out.to(out_dtype)
    T = np.diff(cu).astype(np.int64)
    fake = np.asarray([fake_token_len(int(t), True) for t in T], np.int64)
    res = ([out[cu[b]:cu[b + 1]] for b in range(len(T))], T, fake)
    return res + ([feats[cu[b]:cu[b + 1]] for b in range(len(T))],) if return_features else res


@torch.no_grad()
def inference_batched(model, data_in: Sequence, *, frontend: WavFrontend, tower: AudioTower | None = None,
                      loader: Callable | None = None, key: Sequence[str] | None = None, embed_dtype=None, **kwargs):
    """``[model.inference(data_in=[x], ...) for x in data_in]`` with the audio path batched.

    ``model`` is a built (and ``registry.patch``-ed) ``FunASRNano``; ``kwargs`` are passed through to ``model.inference``
    (tokenizer, device, llm kwargs ...); ``batch_size`` is consumed here.  Items of ``data_in`` are waveforms (Tensor /
    ndarray at ``frontend.fs``) or, with ``loader``, anything the loader turns into one.  Returns the list of
    ``model.inference`` results in input order."""
    if len(data_in) == 0:
        raise ValueError("empty batch")
    tower = tower if tower is not None else getattr(model, "audio_tower_b200", None)
    if tower is None:
        raise ValueError("model has no CUDA audio tower: call funasr_b200.patch(model, kwargs) first or pass tower=")
    kwargs.pop("batch_size", None)
    wavs = [_as_waveform(x, loader) for x in data_in]
    embeds, T, _, feats = audio_embeddings(wavs, frontend, tower, return_features=True)
    results = []
    for b, (w, e) in enumerate(zip(wavs, embeds)):
        # the model's own extract_fbank call (model.py:372) gets this utterance's features back instead of recomputing them
        fe_b = KnownFeatures(frontend, int(w.numel()), feats[b].unsqueeze(0), torch.tensor([int(T[b])], dtype=torch.int64))
        emb = e.unsqueeze(0)  # [1, T_b, llm_dim], what model.audio_adaptor would have returned for this utterance
        if embed_dtype is not None:
            emb = emb.to(embed_dtype)
        lens = torch.tensor([int(T[b])], dtype=torch.int32, device=emb.device)
        k = {"key": [key[b]]} if key is not None else {}
        results.append(model.inference(data_in=[w], frontend=fe_b, audio_embedding=emb, audio_embedding_lens=lens,
                                       **k, **kwargs))
    return results
