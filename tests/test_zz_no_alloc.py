"""``fa_op_ln_linear_static`` (include/funasr_b200.h: the allocation-free LayerNorm -> Linear, scratch in a file-scope
``static __device__`` array) against the existing entry points ``fa_op_layernorm`` + ``fa_op_gemm`` with a caller-owned
intermediate.  No counterpart in the reference.

Collected last on purpose: the file name sorts after every other test file.  It
passes alone in a fresh process (first use of the per-device symbol lookup) and after the other tests (warm library)."""
import ctypes as C

import numpy as np
import pytest
import torch


def _problem(M, N, D, seed):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(M, D, generator=g) * 2.0 + 0.5).cuda()
    gamma = (1.0 + 0.1 * torch.randn(D, generator=g)).cuda()
    beta = (0.1 * torch.randn(D, generator=g)).cuda()
    w = (torch.randn(N, D, generator=g) / D ** 0.5).to(torch.bfloat16).cuda()
    bias = (0.1 * torch.randn(N, generator=g)).cuda()
    return x, gamma, beta, w, bias


def _static(x, gamma, beta, w, bias, out, relu=False):
    from funasr_b200 import _lib

    M, D = x.shape
    N = w.shape[0]
    rc = _lib.load().fa_op_ln_linear_static(x.data_ptr(), gamma.data_ptr(), beta.data_ptr(), w.data_ptr(), w.stride(0),
                                            bias.data_ptr(), out.data_ptr(), out.stride(0), M, N, D, int(relu),
                                            int(out.dtype == torch.float32), _lib.stream_ptr())
    return rc


def _existing(x, gamma, beta, w, bias, out_f32, relu=False):
    from gpu_util import op_gemm, op_layernorm

    rows, _ = op_layernorm(x, gamma, beta, want_bf16=True, want_f32=False)  # caller-owned intermediate
    return op_gemm(rows, w, bias=bias, relu=relu, out_f32=out_f32), rows


@pytest.mark.gpu
def test_static_scratch_variant_matches_existing_entry_points_on_two_calls_in_a_row():
    from funasr_b200 import _lib
    from gpu_util import rel_rms

    lib = _lib.load()
    assert lib.fa_static_ln_capacity() == 2 * 1024 * 1024
    # two different problems back to back: the second is SMALLER than the first, so the shared array still holds rows of
    # the first call beyond the second call's M - they must not leak into its result
    cases = [(3000, 2048, 512, False, 11), (777, 1536, 512, True, 12)]  # (M, N, D, f32 output, seed); ReLU on the bf16 one
    outs, wants = [], []
    for M, N, D, out_f32, seed in cases:  # pre-allocate everything: nothing is allocated between or inside the two calls
        p = _problem(M, N, D, seed)
        want, rows = _existing(*p, out_f32=out_f32, relu=not out_f32)
        outs.append((p, torch.full((M, N), float("nan"), dtype=torch.float32 if out_f32 else torch.bfloat16, device="cuda")))
        wants.append((want, rows))
    # set-up, once per device: the first call looks the array's address up (which also makes the driver load the module
    # image that holds the array); a serving loop does this before it starts
    tiny = _problem(8, 64, 512, 3)
    _lib.check(_static(*tiny, torch.empty(8, 64, dtype=torch.bfloat16, device="cuda")), "fa_op_ln_linear_static")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    free_before, _ = torch.cuda.mem_get_info()
    for (p, out), (_, _), (M, N, D, out_f32, _) in zip(outs, wants, cases):
        _lib.check(_static(*p, out, relu=not out_f32), "fa_op_ln_linear_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before  # torch's allocator was not touched
    free_after, _ = torch.cuda.mem_get_info()
    # ... and nothing was cudaMalloc'ed behind its back (driver view of the device; the slack is smaller than either call's
    # intermediate, 3.1 MB and 0.8 MB, would be)
    assert free_after >= free_before - 512 * 1024
    for (p, out), (want, rows), (M, N, D, out_f32, _) in zip(outs, wants, cases):
        assert not torch.isnan(out.float()).any()
        # the same two kernels on the same values: identical bits; the stated tolerance is the GEMM tests' (2e-2 rel-rms
        # against fp32 is their bar; kernel against kernel it is exact)
        assert rel_rms(out.float(), want.float()) <= 1e-6
        assert torch.equal(out, want)
        # and against plain fp32 PyTorch of the op, at the tolerance of tests/test_gpu_ops.py
        x, gamma, beta, w, bias = p
        ref = torch.nn.functional.layer_norm(x, (D,), gamma, beta, eps=1e-12) @ w.float().t() + bias
        if not out_f32:
            ref = torch.relu(ref)
        assert rel_rms(out.float(), ref) <= 2e-2


@pytest.mark.gpu
def test_static_scratch_variant_rejects_what_does_not_fit():
    from funasr_b200 import _lib

    lib = _lib.load()
    cap = lib.fa_static_ln_capacity()
    D = 512
    M = cap // D + 1  # one row too many
    x, gamma, beta, w, bias = _problem(8, 64, D, 1)
    big = torch.zeros(M, D, device="cuda")
    out = torch.empty(M, 64, dtype=torch.bfloat16, device="cuda")
    rc = _static(big, gamma, beta, w, bias, out)
    assert rc != 0
    msg = lib.fa_last_error().decode()
    assert "exceeds the static scratch" in msg and str(cap) in msg
    # exactly at the limit is accepted
    ok = torch.zeros(cap // D, D, device="cuda")
    out2 = torch.empty(cap // D, 64, dtype=torch.bfloat16, device="cuda")
    assert _static(ok, gamma, beta, w, bias, out2) == 0
    torch.cuda.synchronize()
    # misaligned feature dimension
    x7 = torch.zeros(4, 508, device="cuda")
    assert _static(x7, gamma[:508], beta[:508], w[:, :508].contiguous(), bias, torch.empty(4, 64, dtype=torch.bfloat16, device="cuda")) != 0
    assert "multiple of 8" in lib.fa_last_error().decode()
