"""funasr_b200 — B200 (sm_100a) audio hot path of Fun-ASR-Nano behind the reference's own seams.

    from funasr_b200 import patch
    m, kwargs = FunASRNano.from_pretrained(model=...)      # reference, unchanged
    patch(m, kwargs)                                       # frontend / audio_encoder / audio_adaptor -> CUDA
    m.inference(data_in=[wav])                             # reference, unchanged

Importing this package never touches ``oracle/`` and never falls back to CPU/eager math.
"""
from . import _lib
from ._lib import FunASRB200Error, LIB_PATH
from .encoder import AudioTower, SenseVoiceEncoderSmall, Transformer
from .batched import audio_embeddings, inference_batched
from .ingest import Resample, pack_pcm, pcm_to_mono
from .frontend import WavFrontend, extract_fbank, fake_token_len, load_cmvn, pack_waveforms
from .pipeline import AudioPipeline, StreamingPipeline
from .registry import install, patch, tables
from .sharding import gather_results, shard_contiguous, shard_lpt
from .splice import splice_embeds
from .vllm_route import install_vllm_route, patch_vllm_model

__all__ = ["AudioPipeline", "StreamingPipeline", "AudioTower", "audio_embeddings", "inference_batched", "FunASRB200Error", "LIB_PATH", "Resample", "pack_pcm", "pcm_to_mono", "SenseVoiceEncoderSmall", "Transformer",
           "WavFrontend", "extract_fbank", "fake_token_len", "gather_results", "install", "load_cmvn", "pack_waveforms",
           "install_vllm_route", "patch", "patch_vllm_model", "shard_contiguous", "shard_lpt", "splice_embeds", "tables"]
