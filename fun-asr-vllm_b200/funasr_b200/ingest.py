"""Audio ingest on the GPU: drop-in for the ``torchaudio.transforms.Resample(sr, 16000)`` the reference applies to every
utterance on the host (``/root/reference/infer.py:320-330``, ``infer_kaldidata.py:454-455``), plus the packed form that
feeds ``AudioPipeline`` directly.  Same constructor arguments, same call, same output length
``ceil(new*L/orig)``; the filter bank is torchaudio's (``TA/functional/functional.py:1305-1402``) built on the host once, the
convolution is ``fa_resample_sinc``.  Only ``sinc_interp_hann`` (the default, the one the reference uses) is implemented."""
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
    def forward_packed(self, wav: torch.Tensor, starts, n_samples, out: torch.Tensor | None = None):
        """``wav`` packed f32 PCM on the GPU, utterance b = ``n_samples[b]`` samples at ``starts[b]`` (host arrays).  Returns
        (packed 16 kHz PCM, starts, lengths) with every output utterance starting on a 4-sample boundary, which is what
        ``WavFrontend.forward_packed`` / ``AudioPipeline`` take."""
        _lib.require_device(wav.device)
        starts, n = np.asarray(starts, np.int64), np.asarray(n_samples, np.int64)
        if n.size == 0 or (n <= 0).any():
            raise ValueError("empty batch or empty utterance")
        if self.orig_freq == self.new_freq:
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
            _lib.check(_lib.load().fa_resample_sinc(wav.data_ptr(), p_in, p_len, k.data_ptr(), self.orig, self.new, self.width,
                                                    self.taps_padded, self.frames_per_tile, out.data_ptr(), p_out, p_tiles, B, int(tile_off[-1]),
                                                    _lib.stream_ptr(wav.device)), "fa_resample_sinc")
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
