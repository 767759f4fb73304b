"""``torch.cat([prefix, speech[:len], suffix])`` per utterance (``/root/reference/infer.py:462-466``; the in-place
splice of ``model.py:513-541`` is the same gather with the placeholder span as target) as ONE kernel."""
from __future__ import annotations

import numpy as np
import torch

from . import _lib


@torch.no_grad()
def splice_embeds(speech: torch.Tensor, cu_seqlens, prefix: torch.Tensor | None, suffix: torch.Tensor | None,
                  take=None, out_dtype=torch.bfloat16):
    """``speech`` packed ``[sum T, D]`` (fp32 or bf16) on the GPU, ``cu_seqlens`` host int32 ``[B+1]``; ``prefix`` /
    ``suffix`` ``[P, D]`` / ``[S, D]`` in ``out_dtype``; ``take[b]`` rows of utterance b are used (default all — pass
    ``fake_token_len`` values for the ``model.py`` path).  Returns (flat ``[sum(P+take_b+S), D]``, host offsets ``[B+1]``);
    ``flat[off[b]:off[b+1]]`` is the ``prompt_embeds`` tensor of utterance b."""
    _lib.require_device(speech.device)
    cu = np.asarray(cu_seqlens, dtype=np.int32)
    B, D = len(cu) - 1, speech.shape[1]
    lens = np.diff(cu)
    tk = lens if take is None else np.minimum(np.maximum(np.asarray(take, dtype=np.int64), 0), lens)
    P = 0 if prefix is None else prefix.shape[0]
    S = 0 if suffix is None else suffix.shape[0]
    rows = P + tk + S
    off = np.zeros(B + 1, np.int64)
    off[1:] = np.cumsum(rows)
    out_bf16 = out_dtype == torch.bfloat16
    for t in (prefix, suffix):
        if t is not None and (t.dtype != out_dtype or not t.is_contiguous() or t.device != speech.device):
            raise ValueError("prefix/suffix must be contiguous, on the speech device and of out_dtype")
    if speech.dtype not in (torch.float32, torch.bfloat16) or out_dtype not in (torch.float32, torch.bfloat16):
        raise ValueError("speech / out dtype must be fp32 or bf16")
    dev = speech.device
    meta = torch.from_numpy(np.concatenate([off.view(np.int32), cu, tk.astype(np.int32)])).pin_memory().to(dev, non_blocking=True)
    p_off, p_cu, p_take = meta.data_ptr(), meta[2 * (B + 1):].data_ptr(), meta[3 * (B + 1):].data_ptr()
    out = torch.empty((int(off[-1]), D), dtype=out_dtype, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.load().fa_splice_embeds(speech.data_ptr(), 1 if speech.dtype == torch.bfloat16 else 0, p_cu, p_take,
                                                _lib.ptr(prefix), P, _lib.ptr(suffix), S, p_off, B, D, out.data_ptr(),
                                                1 if out_bf16 else 0, int(rows.max()), _lib.stream_ptr(dev)), "fa_splice_embeds")
    return out, off
