"""ctypes binding of ``libfunasr_b200.so`` (the C ABI declared in ``include/funasr_b200.h``).

There is NO fallback: if the library is missing or the device is not sm_100 the import of any
compute entry point raises.  PyTorch is used only for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libfunasr_b200.so")
ABI_VERSION = 10
FE_TILE_ROWS = 5  # FA_FE_TILE_ROWS

i32, i64, f32, vp, sz = C.c_int32, C.c_int64, C.c_float, C.c_void_p, C.c_size_t


class FrontendCfg(C.Structure):
    _fields_ = [("sample_rate", i32), ("frame_len", i32), ("frame_shift", i32), ("n_fft", i32), ("n_mels", i32),
                ("lfr_m", i32), ("lfr_n", i32), ("preemph", f32), ("upscale", f32), ("log_floor", f32)]


class ModelCfg(C.Structure):
    _fields_ = [("d_in", i32), ("d_model", i32), ("n_heads", i32), ("d_ff", i32), ("fsmn_kernel", i32),
                ("sanm_shift", i32), ("n_blocks", i32), ("n_tp_blocks", i32), ("ad_downsample", i32),
                ("ad_ffn", i32), ("ad_dim", i32), ("ad_heads", i32), ("ad_layers", i32), ("ad_ff", i32)]


class SanmLayerW(C.Structure):
    _fields_ = [(n, vp) for n in ("ln1_g", "ln1_b", "qkv_w", "qkv_b", "fsmn_w", "out_w", "out_b", "ln2_g", "ln2_b",
                                  "ffn1_w", "ffn1_b", "ffn2_w", "ffn2_b", "qkv_colsum", "ffn1_colsum")]


class AdaptorLayerW(C.Structure):
    _fields_ = [(n, vp) for n in ("ln1_g", "ln1_b", "qkv_w", "qkv_b", "out_w", "out_b", "ln2_g", "ln2_b", "ffn1_w",
                                  "ffn1_b", "ffn2_w", "ffn2_b")]


class ModelW(C.Structure):
    _fields_ = [("layers", C.POINTER(SanmLayerW)), ("after_norm_g", vp), ("after_norm_b", vp), ("tp_norm_g", vp),
                ("tp_norm_b", vp), ("pe_inv_timescales", vp), ("lin1_w", vp), ("lin1_b", vp), ("lin2_w", vp),
                ("lin2_b", vp), ("ad_layers", C.POINTER(AdaptorLayerW))]


# name -> (restype, argtypes); must list every symbol include/funasr_b200.h declares
# (tests/test_abi.py parses the header and checks the two agree).
PROTOTYPES = {
    "fa_abi_version": (C.c_int, []),
    "fa_last_error": (C.c_char_p, []),
    "fa_check_device": (C.c_int, [C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]),
    "fa_frontend_tables_bytes": (sz, []),
    "fa_frontend_tables_build": (C.c_int, [C.POINTER(FrontendCfg), vp, vp, vp]),
    "fa_frontend_forward": (C.c_int, [C.POINTER(FrontendCfg), vp, vp, vp, vp, vp, vp, i32, i32, vp, vp, i64, vp]),
    "fa_frontend_logmel": (C.c_int, [C.POINTER(FrontendCfg), vp, vp, vp, vp, vp, i32, i32, vp, i64, vp]),
    "fa_num_frames": (i32, [i32, i32, i32]),
    "fa_num_lfr": (i32, [i32, i32]),
    "fa_fake_token_len": (i32, [i32, i32]),
    "fa_plan_workspace_bytes": (sz, [C.POINTER(ModelCfg), i32]),
    "fa_plan_create": (C.c_int, [C.POINTER(ModelCfg), C.POINTER(ModelW), vp, sz, i32, C.POINTER(vp)]),
    "fa_plan_destroy": (None, [vp]),
    "fa_plan_set_chain": (C.c_int, [vp, i32]),
    "fa_plan_set_graph_rows": (C.c_int, [vp, i32]),
    "fa_plan_set_streams": (C.c_int, [vp, i32]),
    "fa_encoder_forward": (C.c_int, [vp, vp, i32, vp, vp, i32, i32, i32, vp, i32, vp]),
    "fa_adaptor_forward": (C.c_int, [vp, vp, i32, vp, vp, i32, i32, i32, vp, i32, i32, vp]),
    "fa_encode_adapt": (C.c_int, [vp, vp, i32, vp, vp, i32, i32, i32, vp, i32, vp, i32, i32, vp]),
    "fa_splice_embeds": (C.c_int, [vp, i32, vp, vp, vp, i32, vp, i32, vp, i32, i32, vp, i32, i32, vp]),
    "fa_resample_sinc": (C.c_int, [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, vp, i32, i32, vp]),
    "fa_pcm_to_mono_f32": (C.c_int, [vp, i32, i32, f32, vp, vp, vp, vp, i32, i32, vp]),
    "fa_resample_sinc_pcm": (C.c_int, [vp, i32, i32, f32, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, vp, i32, i32, vp]),
    "fa_profile_enable": (C.c_int, [i32]),
    "fa_profile_collect": (C.c_int, [vp, vp, i32]),
    "fa_op_gemm": (C.c_int, [vp, i64, vp, i64, vp, vp, vp, vp, i64, vp, i64, i32, i32, i32, i32, i32, vp]),
    "fa_op_layernorm": (C.c_int, [vp, vp, vp, vp, vp, i32, i32, vp]),
    "fa_static_ln_capacity": (C.c_size_t, []),
    "fa_op_ln_linear_static": (C.c_int, [vp, vp, vp, vp, i64, vp, vp, i64, i32, i32, i32, i32, i32, vp]),
    "fa_op_embed_layernorm": (C.c_int, [vp, vp, vp, i32, f32, vp, vp, vp, i32, i32, vp]),
    "fa_op_gemm_ln": (C.c_int, [vp, i64, vp, i64, vp, vp, vp, i64, i32, i32, i32, i32, i32, i32, vp, C.POINTER(C.c_int32), i32, vp, vp, i64, vp, i64, vp]),
    "fa_op_fsmn": (C.c_int, [vp, i64, i32, vp, vp, i32, vp, i32, i32, i32, i32, vp]),
    "fa_op_fsmn_bf16": (C.c_int, [vp, i64, i32, vp, vp, i32, vp, i64, i32, i32, i32, i32, vp]),
    "fa_op_attention": (C.c_int, [vp, i64, vp, i32, i32, i32, i32, vp, i64, vp]),
    "fa_op_attention_fsmn": (C.c_int, [vp, i64, vp, i32, i32, i32, i32, vp, i64, vp, i32, i32, vp, i64, vp, i64, vp]),
}

PROFILE_KINDS = ("frontend", "gemm", "attention", "layernorm", "fsmn", "repack", "splice")
_lib = None


class FunASRB200Error(RuntimeError):
    pass


def load():
    """dlopen the library once; raises (never falls back) when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FunASRB200Error(
            f"{LIB_PATH} is missing: build it with `make -C fun-asr-vllm_b200/csrc` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU/PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    got = lib.fa_abi_version()
    if got != ABI_VERSION:
        raise FunASRB200Error(f"libfunasr_b200 ABI {got} != binding ABI {ABI_VERSION}; rebuild the library")
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    """Non-zero C return -> RuntimeError(fa_last_error()) (SURVEY.md §8b 'Errors')."""
    if rc != 0:
        msg = load().fa_last_error().decode("utf-8", "replace")
        raise FunASRB200Error(f"{what or 'libfunasr_b200'} failed (rc={rc}): {msg}")


_device_checked = set()


def require_device(device) -> None:
    """Raise unless ``device`` is a CUDA sm_100 device (B200)."""
    import torch

    device = torch.device(device)
    if device.type != "cuda":
        raise FunASRB200Error(f"funasr_b200 runs on CUDA sm_100a devices only, got {device}; there is no CPU fallback")
    idx = device.index if device.index is not None else torch.cuda.current_device()
    if idx in _device_checked:
        return
    with torch.cuda.device(idx):
        sm, mj, mn = i32(), i32(), i32()
        check(load().fa_check_device(C.byref(sm), C.byref(mj), C.byref(mn)), "fa_check_device")
    _device_checked.add(idx)


def ptr(t) -> int | None:
    """Device/host address of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream_ptr(device=None) -> int:
    import torch

    return torch.cuda.current_stream(device).cuda_stream


def profile_enable(on: bool) -> None:
    check(load().fa_profile_enable(1 if on else 0), "fa_profile_enable")


def profile_collect() -> dict:
    """{kernel class: (device ms, launches)} since the last collect (see fa_profile_collect)."""
    n = len(PROFILE_KINDS)
    ms, cnt = (f32 * n)(), (i32 * n)()
    check(load().fa_profile_collect(ms, cnt, n), "fa_profile_collect")
    return {k: (float(ms[i]), int(cnt[i])) for i, k in enumerate(PROFILE_KINDS)}
