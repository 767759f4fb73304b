"""Audio ingest on the GPU: drop-in for the ``torchaudio.transforms.Resample(sr, 16000)`` the reference applies to every
utterance on the host (``/root/reference/infer.py:320-330``, ``infer_kaldidata.py:454-455``), plus the packed form that
feeds ``AudioPipeline`` directly.  Same constructor arguments, same call, same output length
``ceil(new*L/orig)``; the filter bank is torchaudio's (``TA/functional/functional.py:1305-1402``) built on the host once, the
convolution is ``fa_resample_sinc``.  Only ``sinc_interp_hann`` (the default, the one the reference uses) is implemented.

Raw PCM (``pack_pcm`` -> ``Resample.forward_packed_pcm`` / ``pcm_to_mono``): the collator's ``torch.from_numpy(audio).float()``
(``infer.py:320-324``) and ``librosa.load``'s 16-bit scaling and mono mix (``infer_kaldidata.py:101``) happen inside the same
kernel that stages the resampler's input, so int16 / float64 / multi-channel audio is uploaded as it is and read once."""
from __future__ import annotations

import math

import numpy as np
import torch

from . import _lib


def _sinc_kernel(orig: int, new: int, lowpass_filter_width: int, rolloff: float):
    """TA/functional/functional.py:1351-1402 with dtype=None: float64 table (the phase offsets j/new are float32, as torch
    computes an int64 tensor divided by an int), rounded to float32.  ``orig`` / ``new`` already reduced by their gcd."""
    base = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / base)
    idx = np.arange(-width, width + orig, dtype=np.float64)[None, :] / orig
    phase = (np.arange(0, -new, -1, dtype=np.int64).astype(np.float32) / np.float32(new)).astype(np.float64)[:, None]
    t = np.clip((phase + idx) * base, -lowpass_filter_width, lowpass_filter_width)
    window = np.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t = t * math.pi
    with np.errstate(invalid="ignore", divide="ignore"):
        k = np.where(t == 0, 1.0, np.sin(t) / t)
    return (k * (window * (base / orig))).astype(np.float32), width


PCM_F32, PCM_S16, PCM_F64 = 0, 1, 2  # FA_PCM_*
_PCM_OF = {np.dtype(np.float32): PCM_F32, np.dtype(np.int16): PCM_S16, np.dtype(np.float64): PCM_F64}


def pack_pcm(waves):
    """List of numpy arrays ``[L]`` or ``[L, C]`` (frames x channels, the soundfile layout) sharing dtype (int16 / float32 /
    float64) and channel count -> (pinned host buffer of the raw interleaved samples, FA_PCM_* format, channels,
    frame starts, frame counts).  Every utterance starts on a 16-byte boundary."""
    waves = [np.asarray(w) for w in waves]
    if not waves:
        raise ValueError("empty batch")
    dt, ch = waves[0].dtype, (1 if waves[0].ndim == 1 else waves[0].shape[1])
    if dt not in _PCM_OF:
        raise TypeError(f"PCM dtype {dt} is not one of int16, float32, float64")
    for w in waves:
        if w.dtype != dt or w.ndim not in (1, 2) or (1 if w.ndim == 1 else w.shape[1]) != ch or w.shape[0] == 0:
            raise ValueError("utterances of one batch share dtype and channel count and are not empty")
    n = np.array([w.shape[0] for w in waves], np.int64)
    align = 16 // math.gcd(16, dt.itemsize * ch)  # frames per 16 bytes
    starts = np.zeros(len(n), np.int64)
    starts[1:] = np.cumsum((n[:-1] + align - 1) // align * align)
    total = int(starts[-1] + (n[-1] + align - 1) // align * align)
    host = torch.zeros(total * ch, dtype={PCM_F32: torch.float32, PCM_S16: torch.int16, PCM_F64: torch.float64}[_PCM_OF[dt]]).pin_memory()
    flat = host.numpy()
    for w, s0, k in zip(waves, starts, n):
        flat[s0 * ch:(s0 + k) * ch] = np.ascontiguousarray(w).reshape(-1)
    return host, _PCM_OF[dt], ch, starts, n


def _meta(device, *arrays):
    h = torch.from_numpy(np.concatenate([np.ascontiguousarray(a).view(np.int32).reshape(-1) for a in arrays])).pin_memory()
    return h.to(device, non_blocking=True)


@torch.no_grad()
def pcm_to_mono(raw: torch.Tensor, fmt: int, channels: int, starts, n_frames, scale: float = 1.0):
    """Raw interleaved PCM on the GPU -> (packed mono f32, starts, lengths) at the SAME rate (the sr == 16000 branch of the
    reference's collator): ``fa_pcm_to_mono_f32``."""
    _lib.require_device(raw.device)
    starts, n = np.asarray(starts, np.int64), np.asarray(n_frames, np.int64)
    if n.size == 0 or (n <= 0).any():
        raise ValueError("empty batch or empty utterance")
    B = len(n)
    ostarts = np.zeros(B, np.int64)
    ostarts[1:] = np.cumsum((n[:-1] + 3) // 4 * 4)
    total = int(ostarts[-1] + (n[-1] + 3) // 4 * 4)
    meta = _meta(raw.device, starts, ostarts, n.astype(np.int32))
    out = torch.zeros(total, dtype=torch.float32, device=raw.device)
    with torch.cuda.device(raw.device):
        _lib.check(_lib.load().fa_pcm_to_mono_f32(raw.data_ptr(), fmt, channels, float(scale), meta.data_ptr(), meta[4 * B:].data_ptr(),
                                                  out.data_ptr(), meta[2 * B:].data_ptr(), B, int(n.max()), _lib.stream_ptr(raw.device)),
                   "fa_pcm_to_mono_f32")
    out._fa_keep = meta
    return out, ostarts, n


class Resample(torch.nn.Module):
    """``Resample(orig_freq, new_freq)(waveform)``: waveform ``[L]`` or ``[..., L]`` floating point, CPU or CUDA; the result is
    on ``device`` (default: the waveform's device if CUDA, else cuda:0)."""

    def __init__(self, orig_freq: int = 16000, new_freq: int = 16000, resampling_method: str = "sinc_interp_hann",
                 lowpass_filter_width: int = 6, rolloff: float = 0.99, beta=None, *, dtype=None, device=None):
        super().__init__()
        if resampling_method not in ("sinc_interp_hann", "sinc_interpolation"):
            raise ValueError(f"Invalid resampling method: {resampling_method} (only sinc_interp_hann is implemented)")
        if not (int(orig_freq) == orig_freq and int(new_freq) == new_freq):
            raise Exception("Frequencies must be of integer type to ensure quality resampling computation.")
        if lowpass_filter_width <= 0:
            raise ValueError("Low pass filter width should be positive.")
        self.orig_freq, self.new_freq = int(orig_freq), int(new_freq)
        self.gcd = math.gcd(self.orig_freq, self.new_freq)
        self.orig, self.new = self.orig_freq // self.gcd, self.new_freq // self.gcd
        self.lowpass_filter_width, self.rolloff = lowpass_filter_width, rolloff
        self.device = torch.device(device) if device is not None else None
        self.width, self._table, self._table_dev = 0, None, {}
        if self.orig_freq != self.new_freq:
            self._table, self.width = _sinc_kernel(self.orig, self.new, lowpass_filter_width, rolloff)
        # frames per CTA: keep the staged span and the output tile around 48 KB each
        J = 8 if self.new >= 8 else 4 if self.new >= 4 else 2 if self.new >= 2 else 1  # phases per lane (ingest.cu)
        F = 8 // J                                                                    # frames per lane
        fw = max(1, min(64, 12288 // (32 * max(self.orig, self.new))))  # up to 2048 frames per CTA for small ratios
        self.frames_per_tile = 32 * max(F, fw // F * F)
        self.taps_padded = (2 * self.width + self.orig + 3) // 4 * 4

    def out_lengths(self, n_samples) -> np.ndarray:
        n = np.asarray(n_samples, dtype=np.int64)
        return -(-self.new * n // self.orig)

    def _kernel_on(self, device):
        key = str(device)
        if key not in self._table_dev:
            padded = np.zeros((self.new, self.taps_padded), np.float32)  # rows padded to 16-byte multiples with zero taps
            padded[:, : self._table.shape[1]] = self._table
            self._table_dev[key] = torch.from_numpy(padded).to(device)
        return self._table_dev[key]

    @torch.no_grad()
    def forward_packed_pcm(self, raw: torch.Tensor, fmt: int, channels: int, starts, n_frames, scale: float = 1.0):
        """Raw interleaved PCM on the GPU (``pack_pcm``) -> packed mono f32 at ``new_freq``: float conversion, scaling, mono
        mix and resampling in ONE pass (``fa_resample_sinc_pcm``); equal rates take ``pcm_to_mono``."""
        if self.orig_freq == self.new_freq:
            return pcm_to_mono(raw, fmt, channels, starts, n_frames, scale)
        return self.forward_packed(raw, starts, n_frames, pcm=(int(fmt), int(channels), float(scale)))

    @torch.no_grad()
    def forward_packed(self, wav: torch.Tensor, starts, n_samples, out: torch.Tensor | None = None, pcm=(PCM_F32, 1, 1.0)):
        """``wav`` packed f32 PCM on the GPU, utterance b = ``n_samples[b]`` samples at ``starts[b]`` (host arrays).  Returns
        (packed 16 kHz PCM, starts, lengths) with every output utterance starting on a 4-sample boundary, which is what
        ``WavFrontend.forward_packed`` / ``AudioPipeline`` take."""
        _lib.require_device(wav.device)
        starts, n = np.asarray(starts, np.int64), np.asarray(n_samples, np.int64)
        if n.size == 0 or (n <= 0).any():
            raise ValueError("empty batch or empty utterance")
        if self.orig_freq == self.new_freq:
            if pcm != (PCM_F32, 1, 1.0):
                return pcm_to_mono(wav, pcm[0], pcm[1], starts, n, pcm[2])
            return wav, starts, n
        B = len(n)
        on = self.out_lengths(n)
        ostarts = np.zeros(B, np.int64)
        ostarts[1:] = np.cumsum((on[:-1] + 3) // 4 * 4)
        total = int(ostarts[-1] + (on[-1] + 3) // 4 * 4)
        tiles = -(-(n // self.orig + 1) // self.frames_per_tile)
        tile_off = np.zeros(B + 1, np.int32)
        tile_off[1:] = np.cumsum(tiles)
        meta_h = torch.from_numpy(np.concatenate([starts.view(np.int32), ostarts.view(np.int32), n.astype(np.int32), tile_off]))
        meta = meta_h.pin_memory().to(wav.device, non_blocking=True)
        p_in, p_out = meta.data_ptr(), meta[2 * B:].data_ptr()
        p_len, p_tiles = meta[4 * B:].data_ptr(), meta[5 * B:].data_ptr()
        if out is None:
            out = torch.zeros(total, dtype=torch.float32, device=wav.device)
        k = self._kernel_on(wav.device)
        with torch.cuda.device(wav.device):
            _lib.check(_lib.load().fa_resample_sinc_pcm(wav.data_ptr(), pcm[0], pcm[1], pcm[2], p_in, p_len, k.data_ptr(), self.orig,
                                                        self.new, self.width, self.taps_padded, self.frames_per_tile, out.data_ptr(), p_out,
                                                        p_tiles, B, int(tile_off[-1]), _lib.stream_ptr(wav.device)), "fa_resample_sinc_pcm")
        self._keep = meta
        return out, ostarts, on

    @torch.no_grad()
    def forward(self, waveform: torch.Tensor) -> torch.Tensor:
        if not waveform.is_floating_point():
            raise TypeError(f"Expected floating point type for waveform tensor, but received {waveform.dtype}.")
        if self.orig_freq == self.new_freq:
            return waveform
        dev = self.device or (waveform.device if waveform.is_cuda else torch.device("cuda", torch.cuda.current_device()))
        shape = waveform.shape
        x = waveform.reshape(-1, shape[-1]).to(device=dev, dtype=torch.float32).contiguous()
        R, L = x.shape
        if L % 4:  # rows of the packed form start on 16-byte boundaries
            x = torch.nn.functional.pad(x, (0, 4 - L % 4))
        stride = x.shape[1]
        out, ostarts, on = self.forward_packed(x.reshape(-1), np.arange(R, dtype=np.int64) * stride, np.full(R, L, np.int64))
        ol, ostride = int(on[0]), int((on[0] + 3) // 4 * 4)
        y = out.reshape(R, ostride)[:, :ol]
        return y.reshape(*shape[:-1], ol).to(waveform.dtype)
