"""CPU oracle (plain PyTorch, fp32) for the SenseVoice SAN-M encoder and the Transformer adaptor.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The reference looks these classes up by name in funasr's registry
(``/root/reference/model.py:57-59`` encoder, ``:112-118`` adaptor) and calls them at
``/root/reference/infer.py:455-460`` / ``model.py:251-263,488-493``.  funasr is not on this box;
the arithmetic restated here follows the only on-box copy,

  VM = site-packages/vllm/model_executor/models/funasr.py   (vLLM 0.22.0)

with ``ReplicatedLinear / ColumnParallelLinear / RowParallelLinear / QKVParallelLinear`` replaced
by ``nn.Linear`` and the parameter names of the funasr checkpoint (``linear_out``,
``linear_q/k/v``: VM:845-856 maps them to vLLM's names).

Definition of the oracle: **per-utterance, unpadded, batch-of-1 execution**.
On valid rows that equals the reference's padded + key-masked batch (masking is key-only,
the FSMN input is masked, LN/FFN are per token), and it is the one definition on which upstream
funasr (masked adaptor blocks) and vLLM's restatement (unmasked, VM:450) agree.
``forward_padded`` is the literal padded/masked restatement used to prove that equivalence.

Pinned by ``tests/golden/encoder_*.npz`` (outputs of vLLM's classes built on CPU by
``tests/golden/make_golden.py``) in ``tests/test_oracle_encoder.py``.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class LayerNorm(nn.LayerNorm):
    """VM:83-91: eps = 1e-12."""

    def __init__(self, n):
        super().__init__(n, eps=1e-12)


def sinusoidal_pe(T: int, depth: int, dtype=torch.float32) -> torch.Tensor:
    """VM:256-290 ``SinusoidalPositionEncoder``: positions 1..T, ``inv = exp(-k*ln(1e4)/(depth/2-1))``,
    ``[sin | cos]`` concatenated, everything in ``dtype``."""
    pos = torch.arange(1, T + 1).type(dtype)
    inc = torch.log(torch.tensor([10000], dtype=dtype)) / (depth / 2 - 1)
    inv = torch.exp(torch.arange(depth / 2).type(dtype) * (-inc))
    st = pos[:, None] * inv[None, :]
    return torch.cat([torch.sin(st), torch.cos(st)], dim=1).type(dtype)


class SanmAttention(nn.Module):
    """VM:145-253 ``MultiHeadedAttentionSANM``."""

    def __init__(self, n_head, in_feat, n_feat, kernel_size, sanm_shift=0):
        super().__init__()
        self.h, self.d_k = n_head, n_feat // n_head
        self.linear_out = nn.Linear(n_feat, n_feat)
        self.linear_q_k_v = nn.Linear(in_feat, n_feat * 3)
        self.fsmn_block = nn.Conv1d(n_feat, n_feat, kernel_size, groups=n_feat, bias=False)
        left = (kernel_size - 1) // 2 + (sanm_shift if sanm_shift > 0 else 0)  # VM:175-179
        self.pad = (left, kernel_size - 1 - left)

    def fsmn(self, v, mask):
        # VM:181-201; v [B,T,D], mask [B,T,1] or None
        if mask is not None:
            v = v * mask
        x = F.pad(v.transpose(1, 2), self.pad)
        x = self.fsmn_block(x).transpose(1, 2) + v
        return x * mask if mask is not None else x

    def forward(self, x, mask=None):
        # x [B,T,in]; mask [B,T] float (1 valid) or None
        B, T, _ = x.shape
        q, k, v = torch.split(self.linear_q_k_v(x), self.h * self.d_k, dim=-1)  # VM:206-207
        mem = self.fsmn(v, None if mask is None else mask[:, :, None])  # VM:249
        qh = q.reshape(B, T, self.h, self.d_k).transpose(1, 2) * self.d_k ** (-0.5)  # VM:250
        kh = k.reshape(B, T, self.h, self.d_k).transpose(1, 2)
        vh = v.reshape(B, T, self.h, self.d_k).transpose(1, 2)
        scores = qh @ kh.transpose(-2, -1)  # VM:251
        if mask is not None:  # VM:222-230
            km = mask[:, None, None, :].eq(0)
            attn = torch.softmax(scores.masked_fill(km, -float("inf")), dim=-1).masked_fill(km, 0.0)
        else:
            attn = torch.softmax(scores, dim=-1)
        ctx = (attn @ vh).transpose(1, 2).reshape(B, T, self.h * self.d_k)  # VM:235-236
        return self.linear_out(ctx) + mem  # VM:238,253


class FeedForward(nn.Module):
    """VM:412-431."""

    def __init__(self, idim, hidden):
        super().__init__()
        self.w_1 = nn.Linear(idim, hidden)
        self.w_2 = nn.Linear(hidden, idim)

    def forward(self, x):
        return self.w_2(F.relu(self.w_1(x)))


class SanmLayer(nn.Module):
    """VM:94-142 ``EncoderLayerSANM`` (pre-LN; no attention residual when in_size != size)."""

    def __init__(self, in_size, size, n_head, ff, kernel_size, sanm_shift=0):
        super().__init__()
        self.self_attn = SanmAttention(n_head, in_size, size, kernel_size, sanm_shift)
        self.feed_forward = FeedForward(size, ff)
        self.norm1, self.norm2 = LayerNorm(in_size), LayerNorm(size)
        self.in_size, self.size = in_size, size

    def forward(self, x, mask=None):
        a = self.self_attn(self.norm1(x), mask)
        x = x + a if self.in_size == self.size else a  # VM:123-136
        return x + self.feed_forward(self.norm2(x))  # VM:138-140


class SenseVoiceEncoderSmall(nn.Module):
    """VM:293-409."""

    def __init__(self, input_size, output_size=256, attention_heads=4, linear_units=2048,
                 num_blocks=6, tp_blocks=0, kernel_size=11, sanm_shift=0, **kwargs):
        super().__init__()
        self._output_size = output_size
        mk = lambda i: SanmLayer(i, output_size, attention_heads, linear_units, kernel_size, sanm_shift)
        self.encoders0 = nn.ModuleList([mk(input_size)])
        self.encoders = nn.ModuleList([mk(output_size) for _ in range(num_blocks - 1)])
        self.tp_encoders = nn.ModuleList([mk(output_size) for _ in range(tp_blocks)])
        self.after_norm, self.tp_norm = LayerNorm(output_size), LayerNorm(output_size)

    def output_size(self):
        return self._output_size

    def _run(self, x, mask):
        x = x * self._output_size ** 0.5  # VM:388 (the reference does this in place)
        x = x + sinusoidal_pe(x.shape[1], x.shape[2], x.dtype)[None]  # VM:390
        for l in self.encoders0:
            x = l(x, mask)
        for l in self.encoders:
            x = l(x, mask)
        x = self.after_norm(x)  # VM:400
        for l in self.tp_encoders:
            x = l(x, mask)
        return self.tp_norm(x)  # VM:408

    @torch.inference_mode()
    def forward(self, xs_pad, ilens):
        """Oracle definition: each utterance alone, no padding.  Output zero-padded to Tmax."""
        B, Tmax, _ = xs_pad.shape
        out = xs_pad.new_zeros(B, Tmax, self._output_size)
        for b in range(B):
            t = int(ilens[b])
            out[b, :t] = self._run(xs_pad[b : b + 1, :t], None)[0]
        return out, ilens.to(torch.int32)

    @torch.inference_mode()
    def forward_padded(self, xs_pad, ilens):
        """Literal VM:378-409: padded batch with key masks."""
        mask = (torch.arange(xs_pad.shape[1])[None, :] < ilens[:, None]).to(xs_pad.dtype)  # VM:72-80
        return self._run(xs_pad, mask), mask.sum(1).int()


class AdaptorAttention(nn.Module):
    """VM:458-525 ``FunASRAudioAttention`` = plain MHA, scale d_k^-0.5; funasr names linear_q/k/v/out."""

    def __init__(self, n_head, n_feat):
        super().__init__()
        self.h, self.d_k = n_head, n_feat // n_head
        self.linear_q, self.linear_k = nn.Linear(n_feat, n_feat), nn.Linear(n_feat, n_feat)
        self.linear_v, self.linear_out = nn.Linear(n_feat, n_feat), nn.Linear(n_feat, n_feat)

    def forward(self, x):
        B, T, _ = x.shape
        sp = lambda y: y.reshape(B, T, self.h, self.d_k).transpose(1, 2)
        q, k, v = sp(self.linear_q(x)), sp(self.linear_k(x)), sp(self.linear_v(x))
        attn = torch.softmax((q * self.d_k ** (-0.5)) @ k.transpose(-2, -1), dim=-1)
        return self.linear_out((attn @ v).transpose(1, 2).reshape(B, T, self.h * self.d_k))


class AdaptorLayer(nn.Module):
    """VM:434-455 ``EncoderLayer`` (pre-LN)."""

    def __init__(self, size, n_head, ff):
        super().__init__()
        self.self_attn = AdaptorAttention(n_head, size)
        self.feed_forward = FeedForward(size, ff)
        self.norm1, self.norm2 = LayerNorm(size), LayerNorm(size)

    def forward(self, x):
        x = x + self.self_attn(self.norm1(x))
        return x + self.feed_forward(self.norm2(x))


class Transformer(nn.Module):
    """VM:528-595 adaptor ``Transformer`` (Fun-ASR-Nano: downsample_rate 1, ffn 2048, llm 1024,
    2 blocks of 8 heads with FFN llm_dim//4: VM:625-637, VM:555-572)."""

    def __init__(self, downsample_rate=2, encoder_dim=1280, llm_dim=4096, ffn_dim=2048, **kwargs):
        super().__init__()
        self.k, self.encoder_dim, self.llm_dim = downsample_rate, encoder_dim, llm_dim
        self.linear1 = nn.Linear(encoder_dim * self.k, ffn_dim)
        self.linear2 = nn.Linear(ffn_dim, llm_dim)
        n_layer = kwargs.get("n_layer", 2)
        self.blocks = None
        if n_layer > 0:
            self.blocks = nn.ModuleList(
                [AdaptorLayer(llm_dim, kwargs.get("attention_heads", 8), llm_dim // 4) for _ in range(n_layer)]
            )

    def _run(self, x):
        B, T, D = x.shape
        chunk = (T - 1) // self.k + 1  # VM:578-584
        x = F.pad(x, (0, 0, 0, chunk * self.k - T)).reshape(B, chunk, D * self.k)
        x = self.linear2(F.relu(self.linear1(x)))  # VM:585-587
        if self.blocks is not None:
            for blk in self.blocks:
                x = blk(x)
        return x

    @torch.inference_mode()
    def forward(self, x, ilens):
        """Per-utterance oracle definition; output zero-padded to max(olens)."""
        ilens = torch.as_tensor(ilens)
        olens = (ilens - 1) // self.k + 1  # VM:590
        out = x.new_zeros(x.shape[0], int(olens.max()), self.llm_dim)
        for b in range(x.shape[0]):
            t = int(ilens[b])
            out[b, : int(olens[b])] = self._run(x[b : b + 1, :t])[0]
        return out, olens


# ----------------------------------------------------------------------------- helpers
NANO_ENCODER_CONF = dict(output_size=512, attention_heads=4, linear_units=2048, num_blocks=50,
                         tp_blocks=20, kernel_size=11, sanm_shift=0)
NANO_ADAPTOR_CONF = dict(downsample_rate=1, encoder_dim=512, llm_dim=1024, ffn_dim=2048, n_layer=2,
                         attention_heads=8)


def seeded_init_(module: nn.Module, seed: int, round_bf16: bool = True) -> nn.Module:
    """Deterministic random-init used by tests and bench on BOTH sides:
    ``nn.Linear``/``Conv1d`` weights ~ U(-1/sqrt(fan_in), 1/sqrt(fan_in)) from one seeded
    generator in parameter-name order, LN gamma = 1 + 0.1 u, beta = 0.1 u; every matrix
    optionally rounded to bf16 so the GPU path (bf16 weights) and the oracle (fp32 math) see
    identical values."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in sorted(module.named_parameters()):
            u = torch.rand(p.shape, generator=g) * 2 - 1
            if "norm" in name:
                p.copy_(1.0 + 0.1 * u if name.endswith("weight") else 0.1 * u)
            else:
                fan_in = p[0].numel() if p.dim() > 1 else p.numel()
                if name.endswith("bias"):
                    p.copy_(0.1 * u)
                else:
                    p.copy_(u / math.sqrt(fan_in) * (3.0 ** 0.5 if "fsmn" not in name else 1.0))
            if round_bf16 and p.dim() > 1 and "fsmn" not in name:
                p.copy_(p.to(torch.bfloat16).to(torch.float32))
    return module.eval()
