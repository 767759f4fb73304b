"""Helpers shared by the -m gpu parity tests (thin ctypes callers of the C ABI)."""
import ctypes as C

import numpy as np
import torch

from funasr_b200 import _lib


def lib():
    return _lib.load()


def stream():
    return _lib.stream_ptr()


def dev(x, dtype=None):
    t = torch.as_tensor(x)
    if dtype is not None:
        t = t.to(dtype)
    return t.contiguous().cuda()


def cu_of(lens):
    cu = np.zeros(len(lens) + 1, np.int32)
    cu[1:] = np.cumsum(lens)
    return cu


def op_gemm(a_bf16, w_bf16, bias=None, res1=None, res2=None, relu=False, out_f32=False, out=None, res_bf16=None):
    """``out`` given and passed again as ``res1`` / ``res2`` = in-place accumulate (C += ...).  ``res_bf16``: one more
    residual kept in bf16 (any row stride)."""
    M, K = a_bf16.shape
    N = w_bf16.shape[0]
    c = out if out is not None else torch.empty((M, N), dtype=torch.float32 if out_f32 else torch.bfloat16, device="cuda")
    _lib.check(lib().fa_op_gemm(a_bf16.data_ptr(), a_bf16.stride(0), w_bf16.data_ptr(), w_bf16.stride(0), _lib.ptr(bias),
                                _lib.ptr(res1), _lib.ptr(res2), _lib.ptr(res_bf16), res_bf16.stride(0) if res_bf16 is not None else 0,
                                c.data_ptr(), N, M, N, K, int(relu), int(out_f32), stream()),
               "fa_op_gemm")
    return c


def op_gemm_ln_producer(a_bf16, w_bf16, bias, x, inplace=True, xb=None, stats=None, res_bf16=None):
    """x_new = a w^T + bias + x (+ res_bf16) (f32, written over x when inplace), its bf16 copy and the row partials
    [M, n_slots, 2]."""
    import ctypes
    M, K = a_bf16.shape
    N = w_bf16.shape[0]
    c = x if inplace else torch.empty_like(x)
    xb = torch.empty((M, N), dtype=torch.bfloat16, device="cuda") if xb is None else xb
    stats = torch.zeros((M, 16, 2), dtype=torch.float32, device="cuda") if stats is None else stats
    ns = ctypes.c_int32(0)
    _lib.check(lib().fa_op_gemm_ln(a_bf16.data_ptr(), a_bf16.stride(0), w_bf16.data_ptr(), w_bf16.stride(0), bias.data_ptr(),
                                   x.data_ptr(), c.data_ptr(), N, M, N, K, 0, 1, 2, stats.data_ptr(), ctypes.byref(ns), N, None,
                                   xb.data_ptr(), N, _lib.ptr(res_bf16), res_bf16.stride(0) if res_bf16 is not None else 0, stream()),
               "fa_op_gemm_ln")
    n = ns.value
    return c, xb, stats.view(-1)[: M * n * 2].view(M, n, 2), n


def op_gemm_ln_consumer(xb, w_fold_bf16, bias_fold, colsum, stats, n_slots, relu=False, out=None):
    import ctypes
    M, K = xb.shape
    N = w_fold_bf16.shape[0]
    c = out if out is not None else torch.empty((M, N), dtype=torch.bfloat16, device="cuda")
    ns = ctypes.c_int32(n_slots)
    _lib.check(lib().fa_op_gemm_ln(xb.data_ptr(), xb.stride(0), w_fold_bf16.data_ptr(), w_fold_bf16.stride(0), bias_fold.data_ptr(),
                                   None, c.data_ptr(), N, M, N, K, int(relu), 0, 1, stats.data_ptr(), ctypes.byref(ns), K,
                                   colsum.data_ptr(), None, 0, None, 0, stream()), "fa_op_gemm_ln")
    return c


def op_layernorm(x, g, b, want_bf16=True, want_f32=True):
    M, D = x.shape
    yb = torch.empty((M, D), dtype=torch.bfloat16, device="cuda") if want_bf16 else None
    yf = torch.empty((M, D), dtype=torch.float32, device="cuda") if want_f32 else None
    _lib.check(lib().fa_op_layernorm(x.data_ptr(), g.data_ptr(), b.data_ptr(), _lib.ptr(yb), _lib.ptr(yf), M, D, stream()))
    return yb, yf


def op_embed_layernorm(x, inv_ts, cu, scale, g, b):
    M, D = x.shape
    y = torch.empty((M, D), dtype=torch.bfloat16, device="cuda")
    cu_d = dev(cu)
    _lib.check(lib().fa_op_embed_layernorm(x.data_ptr(), inv_ts.data_ptr(), cu_d.data_ptr(), len(cu) - 1, C.c_float(scale),
                                           g.data_ptr(), b.data_ptr(), y.data_ptr(), M, D, stream()))
    return y


def op_fsmn(qkv_bf16, v_col, w, cu, D, K, left):
    M = qkv_bf16.shape[0]
    mem = torch.empty((M, D), dtype=torch.float32, device="cuda")
    cu_d = dev(cu)
    _lib.check(lib().fa_op_fsmn(qkv_bf16.data_ptr(), qkv_bf16.stride(0), v_col, w.data_ptr(), cu_d.data_ptr(), len(cu) - 1,
                                mem.data_ptr(), M, D, K, left, stream()))
    return mem


def op_fsmn_bf16(qkv_bf16, v_col, w, cu, D, K, left, out=None, cu_d=None):
    """bf16 memory written into the right half of a [M, 2D] buffer (the plan's [ctx | mem] layout)."""
    M = qkv_bf16.shape[0]
    cm = torch.zeros((M, 2 * D), dtype=torch.bfloat16, device="cuda") if out is None else out
    cu_d = dev(cu) if cu_d is None else cu_d
    _lib.check(lib().fa_op_fsmn_bf16(qkv_bf16.data_ptr(), qkv_bf16.stride(0), v_col, w.data_ptr(), cu_d.data_ptr(), len(cu) - 1,
                                     cm.data_ptr() + 2 * D, 2 * D, M, D, K, left, stream()))
    return cm


def op_attention(qkv_bf16, cu, n_heads):
    M = qkv_bf16.shape[0]
    dm = n_heads * 128
    ctx = torch.empty((M, dm), dtype=torch.bfloat16, device="cuda")
    cu_d = dev(cu)
    _lib.check(lib().fa_op_attention(qkv_bf16.data_ptr(), qkv_bf16.stride(0), cu_d.data_ptr(), len(cu) - 1,
                                     int(np.diff(cu).max()), n_heads, M, ctx.data_ptr(), dm, stream()))
    return ctx


def op_attention_fsmn(qkv_bf16, w, cu, n_heads, left, acc=None):
    """ctx|mem [M, 2*dm] bf16 from the fused kernel (the plan's layout: memory beside the context).  ``acc`` (f32 [M, dm]):
    the memory is added to it instead (the plan's form for blocks with a residual out-projection) and mem stays zero."""
    M = qkv_bf16.shape[0]
    dm = n_heads * 128
    cm = torch.zeros((M, 2 * dm), dtype=torch.bfloat16, device="cuda")
    cu_d = dev(cu)
    _lib.check(lib().fa_op_attention_fsmn(qkv_bf16.data_ptr(), qkv_bf16.stride(0), cu_d.data_ptr(), len(cu) - 1,
                                          int(np.diff(cu).max()), n_heads, M, cm.data_ptr(), 2 * dm, w.data_ptr(), w.shape[0], left,
                                          cm.data_ptr() + 2 * dm, 2 * dm, _lib.ptr(acc), acc.stride(0) if acc is not None else 0,
                                          stream()))
    return cm


def rel_rms(got, want):
    got, want = got.double().cpu(), want.double().cpu()
    return float((got - want).pow(2).mean().sqrt() / want.pow(2).mean().sqrt().clamp_min(1e-30))
