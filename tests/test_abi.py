"""CPU: the C-ABI library loads and exports every symbol include/funasr_b200.h declares; the ctypes binding lists the
same set; pure-integer entry points (no GPU needed) agree with the oracle."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "funasr_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fa_[a-z0-9_]+)\s*\(", src)))


def test_header_binding_and_library_agree():
    from funasr_b200 import _lib

    declared = _declared()
    assert len(declared) >= 20
    assert sorted(_lib.PROTOTYPES) == declared
    assert os.path.exists(_lib.LIB_PATH), "build the library first: make -C fun-asr-vllm_b200/csrc"
    nm = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
    exported = sorted(set(re.findall(r" T (fa_[a-z0-9_]+)", nm)))
    assert exported == declared  # nothing missing, nothing extra (-fvisibility=hidden)
    lib = _lib.load()
    assert lib.fa_abi_version() == _lib.ABI_VERSION
    assert lib.fa_frontend_tables_bytes() > 4000


def test_length_arithmetic_matches_oracle():
    from funasr_b200 import _lib
    from oracle import frontend as of

    lib = _lib.load()
    for n in [0, 399, 400, 559, 560, 16000, 160000, 160007, 480000]:
        assert lib.fa_num_frames(n, 400, 160) == of.num_frames(n)
    for m in list(range(1, 60)) + [998, 2998]:
        assert lib.fa_num_lfr(m, 6) == of.num_lfr_frames(m)
    for T in list(range(1, 700)):
        for low in (False, True):
            assert lib.fa_fake_token_len(T, int(low)) == of.fake_token_len(T, low)


def test_host_tables_match_torchaudio_golden(frontend_golden):
    import numpy as np

    from funasr_b200.frontend import feature_window, mel_filterbank

    g = frontend_golden
    mb = mel_filterbank()
    assert mb.shape == (80, 257) and np.all(mb[:, 256] == 0)
    np.testing.assert_allclose(mb[:, :256], g["mel_banks"], atol=1e-6)
    for w in ("hamming", "povey"):
        np.testing.assert_allclose(feature_window(w), g[f"window_{w}"], atol=1e-7)


def test_tables_blob_is_sparse_mel(frontend_golden):
    """fa_frontend_tables_build is host-only: the packed blob must hold exactly the non-zero band of each mel row."""
    import ctypes as C

    import numpy as np

    from funasr_b200 import _lib
    from funasr_b200.frontend import FLT_EPS, feature_window, mel_filterbank

    lib = _lib.load()
    cfg = _lib.FrontendCfg(16000, 400, 160, 512, 80, 7, 6, 0.97, 32768.0, FLT_EPS)
    blob = np.zeros(lib.fa_frontend_tables_bytes(), np.uint8)
    win, mel = feature_window("povey"), mel_filterbank()
    _lib.check(lib.fa_frontend_tables_build(C.byref(cfg), win.ctypes.data, mel.ctypes.data, blob.ctypes.data))
    f = blob.view(np.float32)
    np.testing.assert_array_equal(f[:400], win)
    # fp64 constants the lanes build their twiddles from: tw256[t] = W256^t = (cos, -sin)(2 pi t / 256), then
    # sc512[t] = (sin, cos)(2 pi t / 512), t = 0..15
    d = blob[1600:1600 + 2 * 16 * 16].view(np.float64).reshape(2, 16, 2)
    t = np.arange(16)
    np.testing.assert_allclose(d[0, :, 0], np.cos(2 * np.pi * t / 256), atol=1e-15)
    np.testing.assert_allclose(d[0, :, 1], -np.sin(2 * np.pi * t / 256), atol=1e-15)
    np.testing.assert_allclose(d[1, :, 0], np.sin(2 * np.pi * t / 512), atol=1e-15)
    np.testing.assert_allclose(d[1, :, 1], np.cos(2 * np.pi * t / 512), atol=1e-15)
    # sparse mel: first bin of each band (a multiple of four), four-tap steps per group of 16 bands, weights [group][step][lane][4]
    i32 = blob[2112:2112 + 88 * 4].view(np.int32)
    start, steps = i32[:80], i32[80:88]
    assert steps[:5].tolist() == [2, 2, 3, 3, 5] and steps[7] == 1  # the standard bank: exactly the batched path's step counts
    w = blob[2464:2464 + 5 * 5 * 16 * 4 * 4].view(np.float32).reshape(5, 5, 16, 4)
    dense = np.zeros((80, 257 + 24), np.float32)
    for b_ in range(80):
        assert start[b_] % 4 == 0
        for j in range(5):
            dense[b_, start[b_] + 4 * j:start[b_] + 4 * j + 4] += w[b_ // 16, j, b_ % 16]
    np.testing.assert_array_equal(dense[:, :257], mel)
    assert not dense[:, 257:].any()
    bad = _lib.FrontendCfg(16000, 320, 160, 512, 80, 7, 6, 0.97, 32768.0, FLT_EPS)
    assert lib.fa_frontend_tables_build(C.byref(bad), win.ctypes.data, mel.ctypes.data, blob.ctypes.data) != 0
    assert b"not the compiled one" in lib.fa_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "fun-asr-vllm_b200", "funasr_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), fn


def test_cpu_device_is_refused_not_emulated():
    import torch

    import funasr_b200 as fb

    fe = fb.WavFrontend(lfr_m=7, lfr_n=6, dither=0.0, device="cpu")
    with pytest.raises(fb.FunASRB200Error, match="no CPU fallback"):
        fb.extract_fbank([torch.zeros(1600)], frontend=fe)
