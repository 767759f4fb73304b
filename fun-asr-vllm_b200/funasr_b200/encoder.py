"""Drop-in ``SenseVoiceEncoderSmall`` and adaptor ``Transformer`` backed by libfunasr_b200.

The reference builds these by registry name (``/root/reference/model.py:57-59,112-118``) and calls
them as ``audio_encoder(speech, lens)`` / ``audio_adaptor(enc, lens)`` (``infer.py:455,458``,
``model.py:259-263,491``).  The classes here keep

  * the constructor keywords of funasr's classes (on-box copy: VLLM/model_executor/models/funasr.py:293-376, 528-572),
  * the funasr checkpoint parameter names (``encoders0.0.self_attn.linear_q_k_v.weight`` ... ``blocks.1.self_attn.linear_q.weight``),
    so ``load_state_dict`` of a Fun-ASR-Nano checkpoint works unchanged,
  * the call signature and the padded ``[B, T, D]`` fp32 results,

but the ``nn.Linear`` / ``nn.Conv1d`` / ``nn.LayerNorm`` children are parameter containers only: their
``forward`` is never called.  ``forward`` hands packed rows to ``fa_encoder_forward`` / ``fa_adaptor_forward``
(tcgen05 GEMMs, fused LayerNorm / FSMN / attention kernels).  No CPU or eager fallback exists.
"""
from __future__ import annotations

import ctypes as C
import math
import os

import numpy as np
import torch
import torch.nn as nn

from . import _lib


# ----------------------------------------------------------------------------- parameter containers
class _SanmAttnParams(nn.Module):
    def __init__(self, n_head, in_feat, n_feat, kernel_size):
        super().__init__()
        self.linear_out = nn.Linear(n_feat, n_feat)
        self.linear_q_k_v = nn.Linear(in_feat, n_feat * 3)
        self.fsmn_block = nn.Conv1d(n_feat, n_feat, kernel_size, stride=1, padding=0, groups=n_feat, bias=False)


class _FFNParams(nn.Module):
    def __init__(self, idim, hidden):
        super().__init__()
        self.w_1 = nn.Linear(idim, hidden)
        self.w_2 = nn.Linear(hidden, idim)


class _SanmLayerParams(nn.Module):
    def __init__(self, in_size, size, n_head, ff, kernel_size):
        super().__init__()
        self.self_attn = _SanmAttnParams(n_head, in_size, size, kernel_size)
        self.feed_forward = _FFNParams(size, ff)
        self.norm1 = nn.LayerNorm(in_size, eps=1e-12)
        self.norm2 = nn.LayerNorm(size, eps=1e-12)


class _MHAParams(nn.Module):
    def __init__(self, n_feat):
        super().__init__()
        self.linear_q = nn.Linear(n_feat, n_feat)
        self.linear_k = nn.Linear(n_feat, n_feat)
        self.linear_v = nn.Linear(n_feat, n_feat)
        self.linear_out = nn.Linear(n_feat, n_feat)


class _AdaptorLayerParams(nn.Module):
    def __init__(self, size, ff):
        super().__init__()
        self.self_attn = _MHAParams(size)
        self.feed_forward = _FFNParams(size, ff)
        self.norm1 = nn.LayerNorm(size, eps=1e-12)
        self.norm2 = nn.LayerNorm(size, eps=1e-12)


def _fold_ln(norm, weight, bias):
    """LayerNorm(x) W^T + b  ==  rstd * (x (W*gamma)^T - mean * colsum) + (b + W beta): returns the folded weight (fp32, the
    arena rounds it to bf16), the folded bias and the row sums of the bf16-rounded folded weight (what the GEMM multiplies)."""
    w = weight.detach().float()
    wg = w * norm.weight.detach().float()[None, :]
    b = bias.detach().float() + w @ norm.bias.detach().float()
    return wg, b, wg.to(torch.bfloat16).float().sum(dim=1)


# ----------------------------------------------------------------------------- plan (weights arena + workspace)
class _Arena:
    """Packs parameters into two device buffers: bf16 matrices and fp32 vectors, 256-byte aligned slices."""

    def __init__(self, device):
        self.device = device
        self.mats, self.vecs = [], []
        self.n_bf16 = self.n_f32 = 0

    def mat(self, w: torch.Tensor):
        off = self.n_bf16
        self.mats.append((off, w))
        self.n_bf16 += (w.numel() + 127) // 128 * 128
        return ("m", off)

    def vec(self, v: torch.Tensor):
        off = self.n_f32
        self.vecs.append((off, v))
        self.n_f32 += (v.numel() + 63) // 64 * 64
        return ("v", off)

    def commit(self):
        self.bf16 = torch.zeros(max(self.n_bf16, 128), dtype=torch.bfloat16, device=self.device)
        self.f32 = torch.zeros(max(self.n_f32, 64), dtype=torch.float32, device=self.device)
        for off, w in self.mats:
            self.bf16[off:off + w.numel()].copy_(w.detach().reshape(-1))  # fp32 -> bf16 (round to nearest even)
        for off, v in self.vecs:
            self.f32[off:off + v.numel()].copy_(v.detach().reshape(-1))
        self.mats = self.vecs = None

    def addr(self, h):
        kind, off = h
        return (self.bf16.data_ptr() + 2 * off) if kind == "m" else (self.f32.data_ptr() + 4 * off)


class Plan:
    """One ``fa_plan``: encoder and/or adaptor weights + workspace for up to ``max_tokens`` packed rows.

    The weight arena is a SNAPSHOT of the module parameters taken when the plan is built (bf16-rounded matrices, LayerNorm
    folded into the GEMMs): ``load_state_dict`` / ``.to()`` on the module drop the plan (``_PlannedModule.invalidate``), but
    in-place edits of a parameter tensor are not seen until ``invalidate()`` is called."""

    def __init__(self, encoder: "SenseVoiceEncoderSmall | None", adaptor: "Transformer | None", device, max_tokens: int = 4096):
        _lib.require_device(device)
        self.device = torch.device(device)
        self.encoder, self.adaptor = encoder, adaptor
        self.handle = C.c_void_p(None)
        self.cap = 0
        self._pack()
        self._create(max_tokens)

    def twin(self, max_tokens: int | None = None) -> "Plan":
        """A second ``fa_plan`` over the SAME weight arena with its own workspace, so that two half-batches can be in
        flight on two streams (``pipeline.AudioPipeline.run_pinned``); settings (streams, slab size) are inherited."""
        t = Plan.__new__(Plan)
        t.device, t.encoder, t.adaptor = self.device, self.encoder, self.adaptor
        t.handle, t.cap = C.c_void_p(None), 0
        t.arena, t.cfg, t.w = self.arena, self.cfg, self.w
        t._layers, t._ad_layers = getattr(self, "_layers", None), getattr(self, "_ad_layers", None)
        for k in ("graph_rows", "n_streams", "chain"):
            if getattr(self, k, None) is not None:
                setattr(t, k, getattr(self, k))
        t._create(self.cap if max_tokens is None else max_tokens)
        return t

    def _pack(self):
        enc, ad = self.encoder, self.adaptor
        ar = _Arena(self.device)
        cfg = _lib.ModelCfg()
        cfg.d_in = cfg.d_model = 512
        cfg.n_heads, cfg.d_ff, cfg.fsmn_kernel, cfg.sanm_shift = 4, 2048, 11, 0
        cfg.ad_downsample, cfg.ad_ffn, cfg.ad_dim, cfg.ad_heads, cfg.ad_layers, cfg.ad_ff = 1, 2048, 1024, 8, 0, 256
        h = {}
        if enc is not None:
            layers = list(enc.encoders0) + list(enc.encoders) + list(enc.tp_encoders)
            cfg.d_in, cfg.d_model, cfg.n_heads, cfg.d_ff = enc.input_size, enc._output_size, enc.attention_heads, enc.linear_units
            cfg.fsmn_kernel, cfg.sanm_shift = enc.kernel_size, enc.sanm_shift
            cfg.n_blocks, cfg.n_tp_blocks = 1 + len(enc.encoders), len(enc.tp_encoders)
            h["layers"] = []
            fold = os.environ.get("FA_LN_FOLD", "1") != "0"
            for li, L in enumerate(layers):
                a, f = L.self_attn, L.feed_forward
                d = dict()
                # LayerNorm folded into the GEMM it feeds (fa_sanm_layer_w.qkv_colsum / ffn1_colsum): every layer except
                # those whose input comes out of a LayerNorm kernel (layer 0: embed stage; first tp block: after_norm)
                qkv_w, qkv_b, ffn1_w, ffn1_b = a.linear_q_k_v.weight, a.linear_q_k_v.bias, f.w_1.weight, f.w_1.bias
                if fold and li != 0 and li != cfg.n_blocks:
                    qkv_w, qkv_b, cs = _fold_ln(L.norm1, qkv_w, qkv_b)
                    d["qkv_colsum"] = ar.vec(cs)
                if fold and li != 0:
                    ffn1_w, ffn1_b, cs = _fold_ln(L.norm2, ffn1_w, ffn1_b)
                    d["ffn1_colsum"] = ar.vec(cs)
                h["layers"].append(dict(
                    d, ln1_g=ar.vec(L.norm1.weight), ln1_b=ar.vec(L.norm1.bias),
                    qkv_w=ar.mat(qkv_w), qkv_b=ar.vec(qkv_b),
                    fsmn_w=ar.vec(a.fsmn_block.weight.squeeze(1).t().contiguous()),  # [D,1,K] -> tap-major [K,D] f32
                    out_w=ar.mat(a.linear_out.weight), out_b=ar.vec(a.linear_out.bias),
                    ln2_g=ar.vec(L.norm2.weight), ln2_b=ar.vec(L.norm2.bias),
                    ffn1_w=ar.mat(ffn1_w), ffn1_b=ar.vec(ffn1_b),
                    ffn2_w=ar.mat(f.w_2.weight), ffn2_b=ar.vec(f.w_2.bias)))
            h["after_norm_g"], h["after_norm_b"] = ar.vec(enc.after_norm.weight), ar.vec(enc.after_norm.bias)
            h["tp_norm_g"], h["tp_norm_b"] = ar.vec(enc.tp_norm.weight), ar.vec(enc.tp_norm.bias)
            half = enc.input_size // 2
            # SinusoidalPositionEncoder (VLLM/...:268-276): inv = exp(-k * ln(1e4)/(depth/2 - 1)), computed in fp32
            inc = torch.log(torch.tensor([10000.0], dtype=torch.float32)) / (half - 1)
            h["pe"] = ar.vec(torch.exp(torch.arange(half, dtype=torch.float32) * (-inc)))
        if ad is not None:
            if enc is not None and ad.encoder_dim != enc._output_size:
                raise ValueError(f"adaptor encoder_dim {ad.encoder_dim} != encoder output size {enc._output_size}")
            if enc is None:
                cfg.d_model, cfg.n_heads = ad.encoder_dim, ad.encoder_dim // 128
                cfg.d_in = ad.encoder_dim
            cfg.ad_downsample, cfg.ad_ffn, cfg.ad_dim = ad.k, ad.ffn_dim, ad.llm_dim
            blocks = list(ad.blocks) if ad.blocks is not None else []
            cfg.ad_layers, cfg.ad_heads, cfg.ad_ff = len(blocks), ad.attention_heads, ad.llm_dim // 4
            h["lin1_w"], h["lin1_b"] = ar.mat(ad.linear1.weight), ar.vec(ad.linear1.bias)
            h["lin2_w"], h["lin2_b"] = ar.mat(ad.linear2.weight), ar.vec(ad.linear2.bias)
            h["ad_layers"] = []
            for L in blocks:
                a, f = L.self_attn, L.feed_forward
                qkv_w = torch.cat([a.linear_q.weight, a.linear_k.weight, a.linear_v.weight], dim=0)
                qkv_b = torch.cat([a.linear_q.bias, a.linear_k.bias, a.linear_v.bias], dim=0)
                h["ad_layers"].append(dict(
                    ln1_g=ar.vec(L.norm1.weight), ln1_b=ar.vec(L.norm1.bias), qkv_w=ar.mat(qkv_w), qkv_b=ar.vec(qkv_b),
                    out_w=ar.mat(a.linear_out.weight), out_b=ar.vec(a.linear_out.bias),
                    ln2_g=ar.vec(L.norm2.weight), ln2_b=ar.vec(L.norm2.bias),
                    ffn1_w=ar.mat(f.w_1.weight), ffn1_b=ar.vec(f.w_1.bias),
                    ffn2_w=ar.mat(f.w_2.weight), ffn2_b=ar.vec(f.w_2.bias)))
        ar.commit()
        w = _lib.ModelW()
        if enc is not None:
            self._layers = (_lib.SanmLayerW * len(h["layers"]))()
            for i, d in enumerate(h["layers"]):
                for k, v in d.items():
                    setattr(self._layers[i], k, ar.addr(v))
            w.layers = self._layers
            w.after_norm_g, w.after_norm_b = ar.addr(h["after_norm_g"]), ar.addr(h["after_norm_b"])
            w.tp_norm_g, w.tp_norm_b = ar.addr(h["tp_norm_g"]), ar.addr(h["tp_norm_b"])
            w.pe_inv_timescales = ar.addr(h["pe"])
        if ad is not None:
            w.lin1_w, w.lin1_b = ar.addr(h["lin1_w"]), ar.addr(h["lin1_b"])
            w.lin2_w, w.lin2_b = ar.addr(h["lin2_w"]), ar.addr(h["lin2_b"])
            self._ad_layers = (_lib.AdaptorLayerW * max(len(h["ad_layers"]), 1))()
            for i, d in enumerate(h["ad_layers"]):
                for k, v in d.items():
                    setattr(self._ad_layers[i], k, ar.addr(v))
            w.ad_layers = self._ad_layers
        self.arena, self.cfg, self.w = ar, cfg, w

    def _create(self, max_tokens: int):
        lib = _lib.load()
        self.destroy()
        cap = (max(int(max_tokens), 128) + 127) // 128 * 128
        nbytes = int(lib.fa_plan_workspace_bytes(C.byref(self.cfg), cap))
        if nbytes == 0:
            raise _lib.FunASRB200Error("fa_plan_workspace_bytes rejected the model config: " + lib.fa_last_error().decode())
        # zero-filled: rows past the end of a batch must hold finite values (fa_plan_set_graph_rows, include/funasr_b200.h)
        self.workspace = torch.zeros(nbytes + 1024, dtype=torch.uint8, device=self.device)
        base = (self.workspace.data_ptr() + 1023) // 1024 * 1024
        with torch.cuda.device(self.device):
            _lib.check(lib.fa_plan_create(C.byref(self.cfg), C.byref(self.w), base, nbytes, cap, C.byref(self.handle)), "fa_plan_create")
        self.cap = cap
        if getattr(self, "graph_rows", None) is not None:
            _lib.check(lib.fa_plan_set_graph_rows(self.handle, self.graph_rows), "fa_plan_set_graph_rows")
        if getattr(self, "n_streams", None) is not None:
            _lib.check(lib.fa_plan_set_streams(self.handle, self.n_streams), "fa_plan_set_streams")
        if getattr(self, "chain", None) is not None:
            _lib.check(lib.fa_plan_set_chain(self.handle, self.chain), "fa_plan_set_chain")

    def set_graph_rows(self, n: int):
        """Batches of up to ``n`` packed rows replay a captured CUDA graph of the launch program (``fa_plan_set_graph_rows``);
        0 turns replay off."""
        self.graph_rows = int(n)
        _lib.check(_lib.load().fa_plan_set_graph_rows(self.handle, int(n)), "fa_plan_set_graph_rows")

    def set_chain(self, on: bool):
        """True (default): the row-local GEMMs between two attention kernels run as one launch (``fa_plan_set_chain``);
        False: one launch per GEMM.  Bit-identical results."""
        self.chain = int(bool(on))
        _lib.check(_lib.load().fa_plan_set_chain(self.handle, self.chain), "fa_plan_set_chain")

    def set_streams(self, n: int):
        """1 = everything on the caller's stream; 2 = two half-batches on two streams (``fa_plan_set_streams``)."""
        self.n_streams = int(n)
        _lib.check(_lib.load().fa_plan_set_streams(self.handle, int(n)), "fa_plan_set_streams")

    def ensure(self, n_tokens: int):
        if n_tokens > self.cap:
            torch.cuda.current_stream(self.device).synchronize()  # the old workspace may still be in use
            self._create(max(n_tokens, 2 * self.cap))

    def destroy(self):
        if self.handle:
            _lib.load().fa_plan_destroy(self.handle)
            self.handle = C.c_void_p(None)

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    # ---- calls ------------------------------------------------------------------------------
    def _meta(self, lens):
        lens = np.asarray(lens, dtype=np.int64).reshape(-1)
        if lens.size == 0 or (lens <= 0).any():
            raise ValueError("empty batch or zero-length utterance")
        cu = np.zeros(lens.size + 1, np.int32)
        cu[1:] = np.cumsum(lens)
        t = torch.from_numpy(cu)
        return cu, t.pin_memory().to(self.device, non_blocking=True), int(cu[-1]), int(lens.max())

    def encoder_forward(self, feats, feats_tmax, lens, out=None, out_tmax=0, cu_dev=None):
        cu, cu_d, M, tmax = self._meta(lens)
        cu_d = cu_dev if cu_dev is not None else cu_d
        self.ensure(M)
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().fa_encoder_forward(self.handle, feats.data_ptr(), feats_tmax, cu_d.data_ptr(), cu.ctypes.data, len(cu) - 1, M,
                                                      tmax, _lib.ptr(out), out_tmax, _lib.stream_ptr(self.device)),
                       "fa_encoder_forward")
        self._keep = cu_d
        return cu

    def adaptor_forward(self, enc, enc_tmax, lens, out, out_tmax, cu_dev=None):
        cu, cu_d, M, tmax = self._meta(lens)
        cu_d = cu_dev if cu_dev is not None else cu_d
        self.ensure(M)
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().fa_adaptor_forward(self.handle, _lib.ptr(enc), enc_tmax, cu_d.data_ptr(), cu.ctypes.data, len(cu) - 1, M, tmax,
                                                      out.data_ptr(), out_tmax, 1 if out.dtype == torch.bfloat16 else 0,
                                                      _lib.stream_ptr(self.device)), "fa_adaptor_forward")
        self._keep = cu_d
        return cu

    def encode_adapt(self, feats, feats_tmax, lens, out, out_tmax, enc_out=None, enc_out_tmax=0, cu_dev=None):
        cu, cu_d, M, tmax = self._meta(lens)
        cu_d = cu_dev if cu_dev is not None else cu_d
        self.ensure(M)
        with torch.cuda.device(self.device):
            _lib.check(_lib.load().fa_encode_adapt(self.handle, feats.data_ptr(), feats_tmax, cu_d.data_ptr(), cu.ctypes.data, len(cu) - 1, M, tmax,
                                                   _lib.ptr(enc_out), enc_out_tmax, out.data_ptr(), out_tmax,
                                                   1 if out.dtype == torch.bfloat16 else 0, _lib.stream_ptr(self.device)),
                       "fa_encode_adapt")
        self._keep = cu_d
        return cu


def _host_lens(ilens) -> np.ndarray:
    if torch.is_tensor(ilens):
        return ilens.detach().to("cpu").numpy().astype(np.int64)  # one sync if the caller moved lens to the GPU (infer.py:453)
    return np.asarray(ilens, dtype=np.int64)


class _PlannedModule(nn.Module):
    """Common plan cache: (re)built lazily on the device of the parameters, dropped whenever weights change."""

    def __init__(self):
        super().__init__()
        self._plan = None
        self._shared_plan = None  # set by AudioTower: a joint encoder+adaptor plan
        self.register_load_state_dict_post_hook(lambda m, k: m.invalidate())

    def invalidate(self):
        self._plan = None
        if self._shared_plan is not None:
            self._shared_plan.invalidate()

    def _apply(self, fn, *a, **kw):
        self.invalidate()
        return super()._apply(fn, *a, **kw)

    def _device(self):
        return next(self.parameters()).device


# ----------------------------------------------------------------------------- encoder
class SenseVoiceEncoderSmall(_PlannedModule):
    """funasr ``SenseVoiceEncoderSmall`` (VLLM/model_executor/models/funasr.py:293-409) on sm_100a."""

    def __init__(self, input_size: int, output_size: int = 256, attention_heads: int = 4, linear_units: int = 2048,
                 num_blocks: int = 6, tp_blocks: int = 0, dropout_rate: float = 0.1, positional_dropout_rate: float = 0.1,
                 attention_dropout_rate: float = 0.0, stochastic_depth_rate: float = 0.0, input_layer: str | None = "conv2d",
                 pos_enc_class=None, normalize_before: bool = True, concat_after: bool = False,
                 positionwise_layer_type: str = "linear", positionwise_conv_kernel_size: int = 1, padding_idx: int = -1,
                 kernel_size: int = 11, sanm_shfit: int = 0, sanm_shift: int | None = None, selfattention_layer_type: str = "sanm",
                 **kwargs):
        super().__init__()
        if not normalize_before:
            raise ValueError("only normalize_before=True (pre-LN) is compiled")
        if output_size // attention_heads != 128:
            raise ValueError("the attention kernel is compiled for head dim 128 (output_size / attention_heads)")
        self.input_size, self._output_size = input_size, output_size
        self.attention_heads, self.linear_units, self.kernel_size = attention_heads, linear_units, kernel_size
        self.sanm_shift = sanm_shfit if sanm_shift is None else sanm_shift  # funasr spells the keyword "sanm_shfit"
        mk = lambda i: _SanmLayerParams(i, output_size, attention_heads, linear_units, kernel_size)
        self.encoders0 = nn.ModuleList([mk(input_size)])
        self.encoders = nn.ModuleList([mk(output_size) for _ in range(num_blocks - 1)])
        self.tp_encoders = nn.ModuleList([mk(output_size) for _ in range(tp_blocks)])
        self.after_norm = nn.LayerNorm(output_size, eps=1e-12)
        self.tp_norm = nn.LayerNorm(output_size, eps=1e-12)
        for p in self.parameters():
            p.requires_grad_(False)  # inference only (the reference freezes them too, model.py:63-66)

    def output_size(self) -> int:
        return self._output_size

    def plan(self) -> Plan:
        if self._shared_plan is not None:
            return self._shared_plan.plan()
        if self._plan is None:
            self._plan = Plan(self, None, self._device())
        return self._plan

    @torch.no_grad()
    def forward(self, xs_pad: torch.Tensor, ilens, **kwargs):
        """``[B, T, input_size]`` fp32 (zero padded), ``[B]`` -> (``[B, T, output_size]`` fp32 with zero padding rows,
        ``[B]`` int32).  ``xs_pad`` is NOT modified (the reference scales it in place, VLLM/...:388)."""
        plan = self.plan()
        if xs_pad.device != plan.device:
            raise _lib.FunASRB200Error(f"input on {xs_pad.device}, encoder on {plan.device}")
        x = xs_pad.to(torch.float32).contiguous()
        B, T, _ = x.shape
        lens = _host_lens(ilens)
        out = torch.empty((B, T, self._output_size), dtype=torch.float32, device=x.device)
        plan.encoder_forward(x, T, lens, out, T)
        olens = ilens.to(torch.int32) if torch.is_tensor(ilens) else torch.as_tensor(lens, dtype=torch.int32)
        return out, olens


# ----------------------------------------------------------------------------- adaptor
class Transformer(_PlannedModule):
    """funasr adaptor ``Transformer`` (VLLM/model_executor/models/funasr.py:528-595) on sm_100a."""

    def __init__(self, downsample_rate: int = 2, encoder_dim: int = 1280, llm_dim: int = 4096, ffn_dim: int = 2048, **kwargs):
        super().__init__()
        if downsample_rate != 1:
            raise ValueError("only downsample_rate=1 (Fun-ASR-Nano, VLLM/...:628) is compiled")
        self.k, self.encoder_dim, self.llm_dim, self.ffn_dim = downsample_rate, encoder_dim, llm_dim, ffn_dim
        self.attention_heads = kwargs.get("attention_heads", 8)
        if llm_dim // self.attention_heads != 128:
            raise ValueError("the attention kernel is compiled for head dim 128 (llm_dim / attention_heads)")
        self.linear1 = nn.Linear(encoder_dim * self.k, ffn_dim)
        self.linear2 = nn.Linear(ffn_dim, llm_dim)
        n_layer = kwargs.get("n_layer", 2)
        self.blocks = nn.ModuleList([_AdaptorLayerParams(llm_dim, llm_dim // 4) for _ in range(n_layer)]) if n_layer > 0 else None
        for p in self.parameters():
            p.requires_grad_(False)

    def plan(self) -> Plan:
        if self._shared_plan is not None:
            return self._shared_plan.plan()
        if self._plan is None:
            self._plan = Plan(None, self, self._device())
        return self._plan

    @torch.no_grad()
    def forward(self, x: torch.Tensor, ilens=None, **kwargs):
        """``[B, T, encoder_dim]`` fp32, ``[B]`` -> (``[B, T, llm_dim]`` fp32 zero padded, ``[B]`` int32)."""
        plan = self.plan()
        x = x.to(torch.float32).contiguous()
        B, T, _ = x.shape
        lens = np.full(B, T, np.int64) if ilens is None else _host_lens(ilens)
        out = torch.empty((B, T, self.llm_dim), dtype=torch.float32, device=x.device)
        plan.adaptor_forward(x, T, lens, out, T)
        olens = (torch.as_tensor(lens) - 1) // self.k + 1
        return out, olens.to(torch.int32)


# ----------------------------------------------------------------------------- fused tower
class AudioTower:
    """Encoder + adaptor behind ONE plan: ``forward_export(speech, lens)`` of ``/root/reference/model.py:251-254`` without
    materialising the padded fp32 encoder output.  Also the packed fast path used by ``pipeline.AudioPipeline``."""

    def __init__(self, encoder: SenseVoiceEncoderSmall, adaptor: Transformer, max_tokens: int = 4096):
        self.encoder, self.adaptor, self.max_tokens = encoder, adaptor, max_tokens
        self._plan = None
        encoder._shared_plan = self
        adaptor._shared_plan = self

    def invalidate(self):
        self._plan = None

    def plan(self) -> Plan:
        if self._plan is None:
            self._plan = Plan(self.encoder, self.adaptor, self.encoder._device(), self.max_tokens)
        return self._plan

    @torch.no_grad()
    def forward_export(self, speech: torch.Tensor, speech_lengths, out_dtype=torch.float32):
        x = speech.to(torch.float32).contiguous()
        B, T, _ = x.shape
        lens = _host_lens(speech_lengths)
        out = torch.empty((B, T, self.adaptor.llm_dim), dtype=out_dtype, device=x.device)
        self.plan().encode_adapt(x, T, lens, out, T)
        return out, torch.as_tensor(lens, dtype=torch.int32)

    @torch.no_grad()
    def forward_packed(self, feats: torch.Tensor, lens, out: torch.Tensor | None = None, out_dtype=torch.bfloat16, cu_dev=None):
        """Packed ``[sum T, d_in]`` fp32 features -> packed ``[sum T, llm_dim]`` embeddings."""
        lens = _host_lens(lens)
        if out is None:
            out = torch.empty((int(lens.sum()), self.adaptor.llm_dim), dtype=out_dtype, device=feats.device)
        cu = self.plan().encode_adapt(feats, 0, lens, out, 0, cu_dev=cu_dev)
        return out, cu
