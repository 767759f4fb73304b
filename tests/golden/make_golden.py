"""Generate the committed golden fixtures from the REAL third-party code the reference calls.

Run in the build container only (needs torchaudio + vllm importable; neither is read at test
time on the GPU box):

    python tests/golden/make_golden.py

Sources exercised:
  * torchaudio 2.11 ``compliance.kaldi.fbank``                      -> frontend_*.npz
  * vLLM 0.22 ``transformers_utils/processors/funasr.py``: ``WavFrontend`` and
    ``FunASRFeatureExtractor.extract_fbank`` (verbatim copies of funasr's)   -> frontend_*.npz
  * vLLM 0.22 ``model_executor/models/funasr.py``: ``SenseVoiceEncoderSmall`` and the adaptor
    ``Transformer`` built on CPU under a 1-rank gloo TP group               -> encoder_*.npz

Nothing from ``oracle/`` is used to produce the expected outputs (only ``seeded_init_`` to fill
weights deterministically for the full-width case, whose weights are too big to commit).
"""
import os
import sys
import tempfile

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
torch.set_num_threads(1)


def synth(kind: str, n: int, seed: int) -> torch.Tensor:
    g = torch.Generator().manual_seed(seed)
    if kind == "uniform":  # SURVEY.md §8d set 1
        return 0.5 * (2 * torch.rand(n, generator=g) - 1)
    if kind == "tone":  # SURVEY.md §8d set 2
        t = torch.arange(n, dtype=torch.float64) / 16000.0
        return (0.3 * torch.sin(2 * torch.pi * 440.0 * t)).float() + 0.01 * torch.randn(n, generator=g)
    if kind == "zeros":
        return torch.zeros(n)
    raise ValueError(kind)


CMVN_TEXT = None


def cmvn_file(seed=99):
    g = torch.Generator().manual_seed(seed)
    mean = (-(8 + 4 * torch.rand(560, generator=g))).tolist()
    scale = (0.1 + 0.2 * torch.rand(560, generator=g)).tolist()
    txt = "<Nnet>\n<Splice> 560 560\n[ 0 ]\n<AddShift> 560 560\n"
    txt += "<LearnRateCoef> 0 [ " + " ".join(f"{v:.6f}" for v in mean) + " ]\n"
    txt += "<Rescale> 560 560\n"
    txt += "<LearnRateCoef> 0 [ " + " ".join(f"{v:.6f}" for v in scale) + " ]\n</Nnet>\n"
    return txt


def make_frontend():
    import torchaudio.compliance.kaldi as kaldi
    from vllm.transformers_utils.processors.funasr import FunASRFeatureExtractor, WavFrontend

    fx = FunASRFeatureExtractor()
    txt = cmvn_file()
    with tempfile.NamedTemporaryFile("w", suffix=".mvn", delete=False) as f:
        f.write(txt)
        mvn_path = f.name

    out = {"cmvn_text": np.array(txt)}
    # tables
    mb, _ = kaldi.get_mel_banks(80, 512, 16000.0, 20.0, 0.0, 100.0, -500.0, 1.0)
    out["mel_banks"] = mb.numpy()
    for w in ("hamming", "povey"):
        out[f"window_{w}"] = kaldi._feature_window_function(w, 400, 0.42, torch.device("cpu"), torch.float32).numpy()

    # ragged batch through the real extract_fbank/WavFrontend, both windows, cmvn on/off
    specs = [("uniform", 400, 1), ("uniform", 559, 2), ("tone", 560, 3), ("uniform", 16000, 4),
             ("tone", 16007, 5), ("zeros", 4000, 6), ("uniform", 8123, 7), ("tone", 24000, 8)]
    wavs = [synth(k, n, s) for k, n, s in specs]
    out["wav_spec"] = np.array([f"{k}:{n}:{s}" for k, n, s in specs])
    for i, w in enumerate(wavs):
        out[f"wav_{i}"] = w.numpy()
    for win in ("hamming", "povey"):
        for use_cmvn in (False, True):
            fe = WavFrontend(cmvn_file=mvn_path if use_cmvn else None, fs=16000, window=win, n_mels=80,
                             frame_length=25, frame_shift=10, lfr_m=7, lfr_n=6, dither=0.0)
            feats, lens = fx.extract_fbank([w.clone() for w in wavs], frontend=fe, is_final=True)
            tag = f"{win}_{'cmvn' if use_cmvn else 'nocmvn'}"
            out[f"feats_{tag}"] = feats.numpy()
            out[f"lens_{tag}"] = lens.numpy()
    # raw 80-bin log-mel (lfr 1/1) fp32 and fp64 truth for two of them
    for i in (3, 4):
        for win in ("hamming", "povey"):
            for dt, nm in ((torch.float32, "f32"), (torch.float64, "f64")):
                y = kaldi.fbank((wavs[i].to(dt) * 32768)[None], num_mel_bins=80, frame_length=25, frame_shift=10,
                                dither=0.0, energy_floor=0.0, window_type=win, sample_frequency=16000,
                                snip_edges=True)
                out[f"logmel_{i}_{win}_{nm}"] = y.numpy()
    # one 10 s utterance (config 1): keep every 8th LFR row + a checksum to stay small
    w10 = synth("uniform", 160000, 1234)
    for win in ("hamming", "povey"):
        fe = WavFrontend(cmvn_file=None, fs=16000, window=win, n_mels=80, frame_length=25, frame_shift=10,
                         lfr_m=7, lfr_n=6, dither=0.0)
        feats, lens = fx.extract_fbank(w10.clone(), frontend=fe, is_final=True)
        assert tuple(feats.shape) == (1, 167, 560)
        out[f"cfg1_{win}_rows"] = feats[0, ::8].numpy()
        out[f"cfg1_{win}_colsum"] = feats[0].double().sum(0).numpy()
        y64 = kaldi.fbank((w10.double() * 32768)[None], num_mel_bins=80, dither=0.0, energy_floor=0.0,
                          window_type=win)
        out[f"cfg1_{win}_logmel64_rows"] = y64[::48].numpy()
    # LFR reference for many lengths via the real apply_lfr
    from vllm.transformers_utils.processors.funasr import apply_lfr

    ms = list(range(1, 60)) + [97, 98, 998, 1000, 1001, 1002, 1003, 1004, 1005, 2998]
    out["lfr_ms"] = np.array(ms)
    for m in ms:
        mat = torch.arange(m, dtype=torch.float32)[:, None].repeat(1, 2) + torch.tensor([[0.0, 0.5]])
        out[f"lfr_{m}"] = apply_lfr(mat, 7, 6).numpy()[:, ::2].astype(np.int32)  # source frame index per slot
    os.unlink(mvn_path)
    np.savez_compressed(os.path.join(HERE, "frontend_golden.npz"), **out)
    print("frontend_golden.npz", {k: v.shape for k, v in list(out.items())[:6]}, "...")


def vllm_cpu_context():
    from vllm.config import DeviceConfig, VllmConfig, set_current_vllm_config
    from vllm.distributed import parallel_state as ps

    cfg = VllmConfig(device_config=DeviceConfig(device="cpu"))
    ctx = set_current_vllm_config(cfg)
    ctx.__enter__()
    ps.init_distributed_environment(world_size=1, rank=0, distributed_init_method="tcp://127.0.0.1:29541",
                                    local_rank=0, backend="gloo")
    ps.initialize_model_parallel(1, 1)
    return ctx


def to_funasr_names(sd: dict) -> dict:
    """vLLM parameter names -> funasr checkpoint names (inverse of VM:845-856 and VM:641-646)."""
    out = {}
    for k, v in sd.items():
        k2 = k.replace("out_proj.", "linear_out.")
        if ".self_attn.qkv." in k2:
            q, kk, vv = v.chunk(3, dim=0)
            for nm, t in (("linear_q", q), ("linear_k", kk), ("linear_v", vv)):
                out[k2.replace("qkv.", nm + ".")] = t.clone()
            continue
        out[k2] = v.clone()
    return out


def from_funasr_names(sd: dict, adaptor: bool) -> dict:
    out = {}
    for k, v in sd.items():
        out[k.replace("linear_out.", "out_proj.")] = v
    if adaptor:
        res = {}
        for k, v in out.items():
            if ".linear_q." in k:
                base = k.replace("linear_q.", "{}.")
                res[base.format("qkv")] = torch.cat([out[base.format("linear_q")], out[base.format("linear_k")],
                                                      out[base.format("linear_v")]], dim=0)
            elif ".linear_k." in k or ".linear_v." in k:
                continue
            else:
                res[k] = v
        out = res
    return out


def make_encoder():
    ctx = vllm_cpu_context()
    from vllm.model_executor.models import funasr as vf

    from oracle import encoder as oe

    out = {}
    # ---- case A: small dims, weights committed --------------------------------------------
    torch.manual_seed(7)
    small = dict(input_size=48, output_size=64, attention_heads=4, linear_units=96, num_blocks=3, tp_blocks=2,
                 kernel_size=11)
    enc = vf.SenseVoiceEncoderSmall(**small).eval()
    with torch.no_grad():
        for n, p in enc.named_parameters():
            if "norm" in n:
                p.copy_((1.0 if n.endswith("weight") else 0.0) + 0.2 * torch.randn_like(p))
            elif p.dim() == 1:
                p.copy_(0.1 * torch.randn_like(p))
    lens = torch.tensor([20, 13, 5], dtype=torch.int32)
    x = torch.randn(3, 20, 48)
    for b in range(3):
        x[b, lens[b]:] = 0
    with torch.inference_mode():
        y_pad, ol = enc(x.clone(), lens)
        y_single = [enc(x[b:b + 1, :lens[b]].clone(), lens[b:b + 1])[0][0] for b in range(3)]
    out["A_cfg"] = np.array(repr(small))
    out["A_x"], out["A_lens"], out["A_y_pad"], out["A_olens"] = x.numpy(), lens.numpy(), y_pad.numpy(), ol.numpy()
    for b in range(3):
        out[f"A_y_single_{b}"] = y_single[b].numpy()
    for k, v in to_funasr_names(enc.state_dict()).items():
        out["A_w/" + k] = v.numpy()

    ad_small = dict(downsample_rate=1, encoder_dim=64, llm_dim=128, ffn_dim=96, n_layer=2, attention_heads=8)
    ad = vf.Transformer(**ad_small).eval()
    with torch.no_grad():
        for n, p in ad.named_parameters():
            if "norm" in n:
                p.copy_((1.0 if n.endswith("weight") else 0.0) + 0.2 * torch.randn_like(p))
            elif p.dim() == 1:
                p.copy_(0.1 * torch.randn_like(p))
    with torch.inference_mode():
        z_single = [ad(y_single[b][None], lens[b:b + 1])[0][0] for b in range(3)]
    out["A_ad_cfg"] = np.array(repr(ad_small))
    for b in range(3):
        out[f"A_z_single_{b}"] = z_single[b].numpy()
    for k, v in to_funasr_names(ad.state_dict()).items():
        out["A_adw/" + k] = v.numpy()

    # ---- case B: Fun-ASR-Nano widths (560/512/2048/4 heads; adaptor 512->2048->1024, 8 heads), 2+1 blocks,
    #      weights regenerated from the seed at test time ---------------------------------------
    conf = dict(oe.NANO_ENCODER_CONF, num_blocks=2, tp_blocks=1)
    o_enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **conf), seed=4321)
    o_ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)
    v_enc = vf.SenseVoiceEncoderSmall(input_size=560, **conf).eval()
    v_ad = vf.Transformer(**oe.NANO_ADAPTOR_CONF).eval()
    missing = v_enc.load_state_dict(from_funasr_names(o_enc.state_dict(), False), strict=True)
    missing2 = v_ad.load_state_dict(from_funasr_names(o_ad.state_dict(), True), strict=True)
    print("load:", missing, missing2)
    g = torch.Generator().manual_seed(55)
    lensB = torch.tensor([9, 5], dtype=torch.int32)
    xB = torch.randn(2, 9, 560, generator=g)
    xB[1, 5:] = 0
    with torch.inference_mode():
        yB, olB = v_enc(xB.clone(), lensB)
        zB = [v_ad(yB[b:b + 1, :lensB[b]], lensB[b:b + 1])[0][0] for b in range(2)]
    out["B_x"], out["B_lens"], out["B_y_pad"] = xB.numpy(), lensB.numpy(), yB.numpy()
    for b in range(2):
        out[f"B_z_single_{b}"] = zB[b].numpy()
    np.savez_compressed(os.path.join(HERE, "encoder_golden.npz"), **out)
    print("encoder_golden.npz written;", len(out), "arrays")
    ctx.__exit__(None, None, None)


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "frontend"):
        make_frontend()
    if which in ("all", "encoder"):
        make_encoder()
