"""Golden fixture for the vLLM-native multimodal route.

Run in the build container only (needs vllm importable):

    python tests/golden/make_golden_vllm_route.py

Expected outputs come from the REAL ``FunASRForConditionalGeneration.embed_multimodal`` ->
``FunASRModel.get_encoder_outputs`` (VLLM/model_executor/models/funasr.py:967-976, 694-713) called on objects that hold
vLLM's own ``SenseVoiceEncoderSmall`` / ``Transformer`` (CPU, fp32, 1-rank gloo TP group).  The two wrapper classes are
instantiated without their ``__init__`` (which needs a checkpoint's VllmConfig, tokenizer files and the Qwen3 decoder);
the methods under test are the stock ones.  Nothing from ``oracle/`` produces an expected value: only ``seeded_init_``
fills the weights (Fun-ASR-Nano widths, 2+1 blocks + 2 adaptor blocks, regenerated from the seeds at test time).

Also stored: the parameter names and shapes of vLLM's two audio modules at the FULL Fun-ASR-Nano depth (50+20, 2), so
that the name mapping is pinned against the real classes by a CPU test that does not import vllm.
"""
import os
import sys

import numpy as np
import torch
import torch.nn as nn

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, HERE)
torch.set_num_threads(1)

from make_golden import from_funasr_names, vllm_cpu_context  # noqa: E402

LENS = [9, 5, 14]
T_PAD = 20  # the feature extractor pads every item to a fixed max_length (VLLM/transformers_utils/processors/funasr.py:355-362)


def main():
    ctx = vllm_cpu_context()
    from vllm.model_executor.models import funasr as vf
    from vllm.transformers_utils.processors.funasr import FunASRFeatureExtractor  # noqa: F401  (fake_token_len formula below is its)

    from oracle import encoder as oe

    out = {}
    conf = dict(oe.NANO_ENCODER_CONF, num_blocks=2, tp_blocks=1)
    o_enc = oe.seeded_init_(oe.SenseVoiceEncoderSmall(560, **conf), seed=4321)
    o_ad = oe.seeded_init_(oe.Transformer(**oe.NANO_ADAPTOR_CONF), seed=4322)
    v_enc = vf.SenseVoiceEncoderSmall(input_size=560, **conf).eval()
    v_ad = vf.Transformer(**oe.NANO_ADAPTOR_CONF).eval()
    v_enc.load_state_dict(from_funasr_names(o_enc.state_dict(), False), strict=True)
    v_ad.load_state_dict(from_funasr_names(o_ad.state_dict(), True), strict=True)

    holder = nn.Module()
    holder.audio_encoder, holder.audio_adaptor = v_enc, v_ad
    inner = vf.FunASRModel.__new__(vf.FunASRModel)
    nn.Module.__init__(inner)
    inner.encoder = holder
    model = vf.FunASRForConditionalGeneration.__new__(vf.FunASRForConditionalGeneration)
    nn.Module.__init__(model)
    model.model = inner

    g = torch.Generator().manual_seed(77)
    x = torch.randn(len(LENS), T_PAD, 560, generator=g)
    for b, n in enumerate(LENS):
        x[b, n:] = 0
    lens = torch.tensor(LENS, dtype=torch.int32)
    # the feature extractor's own arithmetic (VLLM/transformers_utils/processors/funasr.py:348-352)
    ol = 1 + (lens - 3 + 2 * 1) // 2
    ol = 1 + (ol - 3 + 2 * 1) // 2
    fake = torch.clamp((ol - 1) // 2 + 1, min=1)
    out["x"], out["lens"], out["fake_token_lengths"] = x.numpy(), lens.numpy(), fake.numpy()
    with torch.inference_mode():
        for b in range(len(LENS)):
            y = model.embed_multimodal(input_features=x[b:b + 1].clone(), speech_lengths=lens[b:b + 1],
                                       fake_token_lengths=fake[b:b + 1])
            assert tuple(y.shape) == (1, LENS[b], 1024), y.shape  # the adaptor cuts at max(ilens) (…:575-576)
            out[f"embed_{b}"] = y[0].numpy()

    # names / shapes of the real modules at full depth (meta device: no memory)
    full = dict(oe.NANO_ENCODER_CONF)
    with torch.device("meta"):
        f_enc = vf.SenseVoiceEncoderSmall(input_size=560, **full)
        f_ad = vf.Transformer(**oe.NANO_ADAPTOR_CONF)
    for tag, m in (("enc", f_enc), ("ad", f_ad)):
        sd = m.state_dict()
        out[f"names_{tag}"] = np.array(list(sd.keys()))
        out[f"shapes_{tag}"] = np.array([",".join(str(d) for d in v.shape) for v in sd.values()])
    np.savez_compressed(os.path.join(HERE, "vllm_route_golden.npz"), **out)
    print("vllm_route_golden.npz:", len(out), "arrays;", len(out["names_enc"]), "+", len(out["names_ad"]), "parameter names")
    ctx.__exit__(None, None, None)


if __name__ == "__main__":
    main()
