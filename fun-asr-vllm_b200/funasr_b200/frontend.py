"""Drop-in ``WavFrontend`` / ``extract_fbank`` backed by the fused sm_100a frontend kernel.

Mirrors funasr's ``WavFrontend`` (on-box copy: ``VLLM/transformers_utils/processors/funasr.py:89-168``)
and ``extract_fbank`` (``:292-317``), the functions the reference imports at
``/root/reference/model.py:17`` / ``infer.py:17`` and calls at ``infer.py:446``,
``infer_kaldidata.py:708,897``, ``model.py:372``.  Same constructor keywords, attributes
(``fs, window, n_mels, frame_length, frame_shift, lfr_m, lfr_n, dither, cmvn_file, cmvn``),
``output_size()`` and return types; the arithmetic runs in ONE CUDA kernel over packed offsets
instead of a Python loop over utterances.

Differences, all deliberate and stated in DESIGN.md:
  * the kernel runs dither-off whatever ``dither`` says (dither is random noise added per call, VLLM:123 default 1.0,
    TA:179-181; the path under test is the deterministic one — SURVEY.md §9 Q1); a non-zero ``dither`` is kept as an
    attribute and warned about once;
  * only the 25 ms / 10 ms / 80-mel / LFR 7-6 geometry is compiled;
  * utterances shorter than one frame raise ``ValueError`` (torchaudio asserts, TA:142).
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch

from . import _lib

FLT_EPS = 1.1920928955078125e-07  # TA:22


# ----------------------------------------------------------------------------- host tables
def feature_window(window_type: str, n: int = 400) -> np.ndarray:
    """``_feature_window_function`` (TA:86-113), non-periodic, fp32."""
    k = torch.arange(n, dtype=torch.float32)
    a = 2.0 * math.pi / (n - 1)
    if window_type == "hanning":
        w = 0.5 - 0.5 * torch.cos(a * k)
    elif window_type == "hamming":
        w = 0.54 - 0.46 * torch.cos(a * k)
    elif window_type == "povey":
        w = (0.5 - 0.5 * torch.cos(a * k)).pow(0.85)
    elif window_type == "rectangular":
        w = torch.ones(n)
    else:
        raise ValueError("Invalid window type " + window_type)
    return w.numpy().astype(np.float32)


def mel_filterbank(n_mels: int = 80, n_fft: int = 512, fs: float = 16000.0, low: float = 20.0, high: float = 0.0) -> np.ndarray:
    """``get_mel_banks`` (TA:436-511, no VTLN) + the zero Nyquist column of TA:627 -> [n_mels, n_fft/2+1] fp32."""
    nb = n_fft // 2
    nyq = 0.5 * fs
    if high <= 0.0:
        high += nyq
    mel = lambda f: 1127.0 * math.log(1.0 + f / 700.0)
    mlo, mhi = mel(low), mel(high)
    delta = (mhi - mlo) / (n_mels + 1)
    b = torch.arange(n_mels, dtype=torch.float32).unsqueeze(1)
    left, center, right = mlo + b * delta, mlo + (b + 1.0) * delta, mlo + (b + 2.0) * delta
    freqs = (fs / n_fft) * torch.arange(nb, dtype=torch.float32)
    m = (1127.0 * (1.0 + freqs / 700.0).log()).unsqueeze(0)
    bins = torch.max(torch.zeros(1), torch.min((m - left) / (center - left), (right - m) / (right - center)))
    return torch.nn.functional.pad(bins, (0, 1)).numpy().astype(np.float32)


def load_cmvn(cmvn_file: str) -> torch.Tensor:
    """Kaldi-nnet ``am.mvn`` -> [2, dim] (means; scales).  VLLM:63-86."""
    with open(cmvn_file, encoding="utf-8") as f:
        lines = f.readlines()
    means, scales = [], []
    for i, line in enumerate(lines):
        item = line.split()
        if not item:
            continue
        if item[0] == "<AddShift>":
            nxt = lines[i + 1].split()
            if nxt[0] == "<LearnRateCoef>":
                means = nxt[3:len(nxt) - 1]
        elif item[0] == "<Rescale>":
            nxt = lines[i + 1].split()
            if nxt[0] == "<LearnRateCoef>":
                scales = nxt[3:len(nxt) - 1]
    return torch.as_tensor(np.array([np.array(means, dtype=np.float64), np.array(scales, dtype=np.float64)]),
                           dtype=torch.float32)


def fake_token_len(n_lfr: int, use_low_frame_rate: bool) -> int:
    """``/root/reference/model.py:391-396`` through the C ABI."""
    return int(_lib.load().fa_fake_token_len(int(n_lfr), 1 if use_low_frame_rate else 0))


_dither_warned = False


def _warn_dither(dither) -> None:
    """funasr's default ``dither=1.0`` adds N(0,1) noise to the x32768 samples of every call; the CUDA frontend is
    deterministic and ignores it.  Said once per process, so that an unchanged config.yaml still builds."""
    global _dither_warned
    if not _dither_warned:
        import warnings

        warnings.warn(f"funasr_b200.WavFrontend: dither={dither} is ignored, the CUDA frontend runs dither-off "
                      "(set dither: 0.0 in frontend_conf to silence this)", stacklevel=3)
        _dither_warned = True


# ----------------------------------------------------------------------------- frontend
class WavFrontend(torch.nn.Module):
    """Same surface as funasr's ``WavFrontend`` (VLLM:89-168)."""

    def __init__(self, cmvn_file: str | None = None, fs: int = 16000, window: str = "hamming", n_mels: int = 80,
                 frame_length: int = 25, frame_shift: int = 10, filter_length_min: int = -1,
                 filter_length_max: int = -1, lfr_m: int = 1, lfr_n: int = 1, dither: float = 1.0,
                 snip_edges: bool = True, upsacle_samples: bool = True, device: str | torch.device = "cuda",
                 cmvn: torch.Tensor | None = None, **kwargs):
        super().__init__()
        self.fs, self.window, self.n_mels = fs, window, n_mels
        self.frame_length, self.frame_shift = frame_length, frame_shift
        self.filter_length_min, self.filter_length_max = filter_length_min, filter_length_max
        self.lfr_m, self.lfr_n, self.cmvn_file, self.dither = lfr_m, lfr_n, cmvn_file, dither
        self.snip_edges, self.upsacle_samples = snip_edges, upsacle_samples
        self.cmvn = cmvn if cmvn is not None else (None if cmvn_file is None else load_cmvn(cmvn_file))
        self.device = torch.device(device)
        self._tables = {}  # device index -> (tables blob, cmvn) device tensors
        if dither != 0.0:
            _warn_dither(dither)
        if not snip_edges:
            raise ValueError("only snip_edges=True is compiled")
        self._cfg = _lib.FrontendCfg(fs, int(frame_length * fs / 1000), int(frame_shift * fs / 1000), 512, n_mels,
                                     lfr_m, lfr_n, 0.97, 32768.0 if upsacle_samples else 1.0, FLT_EPS)

    def output_size(self) -> int:
        return self.n_mels * self.lfr_m

    # -- device tables (built on the host once per device through fa_frontend_tables_build)
    def _device_tables(self, device: torch.device):
        key = device.index if device.index is not None else torch.cuda.current_device()
        if key not in self._tables:
            lib = _lib.load()
            blob = torch.empty(int(lib.fa_frontend_tables_bytes()), dtype=torch.uint8)
            win = np.ascontiguousarray(feature_window(self.window, self._cfg.frame_len))
            mel = np.ascontiguousarray(mel_filterbank(self.n_mels, self._cfg.n_fft, float(self.fs)))
            _lib.check(lib.fa_frontend_tables_build(C.byref(self._cfg), win.ctypes.data, mel.ctypes.data, blob.data_ptr()),
                       "fa_frontend_tables_build")
            cm = None
            if self.cmvn is not None:
                cm = torch.as_tensor(self.cmvn, dtype=torch.float32).reshape(2, -1)
                if cm.shape[1] != self.output_size():
                    raise ValueError(f"cmvn dim {cm.shape[1]} != frontend output size {self.output_size()}")
                cm = cm.contiguous().to(device)
            self._tables[key] = (blob.to(device), cm)
        return self._tables[key]

    def lengths(self, n_samples) -> tuple[np.ndarray, np.ndarray]:
        """(fbank frames m, LFR frames T) per utterance: TA:62-66, VLLM:43."""
        n = np.asarray(n_samples, dtype=np.int64)
        m = np.where(n >= self._cfg.frame_len, 1 + (n - self._cfg.frame_len) // self._cfg.frame_shift, 0)
        return m, -(-m // self.lfr_n)

    def forward_packed(self, wav: torch.Tensor, starts: np.ndarray, n_samples: np.ndarray, out: torch.Tensor | None = None,
                       out_tmax: int = 0):
        """Core call.  ``wav`` packed f32 PCM on the GPU; ``starts`` / ``n_samples`` host arrays (first sample and
        sample count of each utterance).  Returns (features, cu_seqlens_host int32 [B+1]); features are packed
        [sum T, 560] (``out_tmax == 0``) or padded [B, out_tmax, 560] with zero padding rows."""
        _lib.require_device(wav.device)
        lib = _lib.load()
        m, T = self.lengths(n_samples)
        if len(T) == 0:
            raise ValueError("empty batch")
        if (m < 1).any():
            bad = int(np.argmin(m))
            raise ValueError(f"utterance {bad} has {int(n_samples[bad])} samples, shorter than one {self.frame_length} ms frame")
        B, D = len(T), self.output_size()
        cu = np.zeros(B + 1, np.int32)
        cu[1:] = np.cumsum(T)
        tiles = -(-T // _lib.FE_TILE_ROWS)
        tile_off = np.zeros(B + 1, np.int32)
        tile_off[1:] = np.cumsum(tiles)
        rows = (np.arange(B, dtype=np.int32) * out_tmax) if out_tmax > 0 else cu[:-1]
        if out_tmax > 0 and out_tmax < int(T.max()):
            raise ValueError("out_tmax smaller than the longest utterance")
        # one small H2D for all per-utterance metadata: int64 starts | int32 lens | rows | tile_off
        st64 = np.ascontiguousarray(starts, dtype=np.int64)
        i32s = np.concatenate([np.asarray(n_samples, np.int32), rows.astype(np.int32), tile_off])
        meta_h = torch.from_numpy(np.concatenate([st64.view(np.int32), i32s]))
        if wav.is_cuda:
            meta_h = meta_h.pin_memory()
        meta = meta_h.to(wav.device, non_blocking=True)
        p_start, p_len = meta.data_ptr(), meta[2 * B:].data_ptr()
        p_rows, p_tiles = meta[3 * B:].data_ptr(), meta[4 * B:].data_ptr()
        if out is None:
            shape = (B, out_tmax, D) if out_tmax > 0 else (int(cu[-1]), D)
            out = (torch.zeros if out_tmax > 0 else torch.empty)(shape, dtype=torch.float32, device=wav.device)
        tab, cm = self._device_tables(wav.device)
        with torch.cuda.device(wav.device):
            _lib.check(lib.fa_frontend_forward(C.byref(self._cfg), tab.data_ptr(), wav.data_ptr(), p_start, p_len,
                                               p_rows, p_tiles, B, int(tile_off[-1]), _lib.ptr(cm),
                                               out.data_ptr(), D, _lib.stream_ptr(wav.device)), "fa_frontend_forward")
        self._keep = meta  # keep the metadata alive until the next call on this stream
        return out, cu

    def logmel(self, wav: torch.Tensor, starts: np.ndarray, n_samples: np.ndarray) -> tuple[torch.Tensor, np.ndarray]:
        """Raw 80-bin log-mel (``kaldi.fbank`` output, before LFR/CMVN), packed [sum m, 80]."""
        _lib.require_device(wav.device)
        m, _ = self.lengths(n_samples)
        B = len(m)
        cu = np.zeros(B + 1, np.int32)
        cu[1:] = np.cumsum(m)
        rows = torch.from_numpy(cu[:-1].copy()).to(wav.device)
        st = torch.from_numpy(np.ascontiguousarray(starts, dtype=np.int64)).to(wav.device)
        ln = torch.from_numpy(np.asarray(n_samples, np.int32)).to(wav.device)
        out = torch.empty((int(cu[-1]), self.n_mels), dtype=torch.float32, device=wav.device)
        tab, _ = self._device_tables(wav.device)
        with torch.cuda.device(wav.device):
            _lib.check(_lib.load().fa_frontend_logmel(C.byref(self._cfg), tab.data_ptr(), wav.data_ptr(), st.data_ptr(),
                                                      ln.data_ptr(), rows.data_ptr(), B, int(m.max()), out.data_ptr(),
                                                      self.n_mels, _lib.stream_ptr(wav.device)), "fa_frontend_logmel")
        return out, cu

    def forward(self, input: torch.Tensor, input_lengths, **kwargs) -> tuple[torch.Tensor, torch.Tensor]:
        """VLLM:128-168: ``input`` [B, Nmax] f32 (zero padded), ``input_lengths`` list/tensor of sample counts ->
        (``[B, Tmax, n_mels*lfr_m]`` f32 zero padded, ``[B]`` lengths).  The result lives on ``self.device``."""
        if input.dim() == 1:
            input = input[None]
        B = input.size(0)
        lens = np.asarray([int(x) for x in (input_lengths.tolist() if torch.is_tensor(input_lengths) else input_lengths)],
                          dtype=np.int64)
        wavs = [input[i, :lens[i]] for i in range(B)]
        wav, starts = pack_waveforms(wavs, self.device)
        _, T = self.lengths(lens)
        feats, cu = self.forward_packed(wav, starts, lens, out_tmax=int(T.max()))
        return feats, torch.from_numpy(np.diff(cu).astype(np.int64))


def pack_waveforms(wavs, device) -> tuple[torch.Tensor, np.ndarray]:
    """list of 1-D f32 tensors (CPU or GPU) -> (packed PCM on ``device``, host int64 start sample of each utterance).
    Each start is rounded up to 4 samples so 16-byte loads stay aligned; CPU inputs are staged through ONE pinned
    buffer and ONE H2D copy."""
    lens = [int(w.numel()) for w in wavs]
    starts = np.zeros(len(lens), np.int64)
    pos = 0
    for i, n in enumerate(lens):
        starts[i] = pos
        pos += (n + 3) // 4 * 4
    total = max(pos, 4)
    if all(w.device.type == "cpu" for w in wavs):
        host = torch.zeros(total, dtype=torch.float32)
        if torch.device(device).type == "cuda":
            host = host.pin_memory()
        for w, s, n in zip(wavs, starts, lens):
            host[s:s + n] = w.reshape(-1)
        wav = host.to(device, non_blocking=True)
    else:
        wav = torch.zeros(total, dtype=torch.float32, device=device)
        for w, s, n in zip(wavs, starts, lens):
            wav[s:s + n] = w.reshape(-1).to(device=device, dtype=torch.float32)
    return wav, starts


def extract_fbank(data, data_len=None, data_type: str = "sound", frontend: WavFrontend | None = None, **kwargs):
    """Same contract as funasr's ``extract_fbank`` (VLLM:292-317): ndarray / Tensor / list thereof ->
    (``[B, Tmax, 560]`` f32, ``[B]`` int32)."""
    if frontend is None:
        raise ValueError("extract_fbank needs frontend=")
    if isinstance(data, np.ndarray):
        data = torch.from_numpy(data)
    if isinstance(data, torch.Tensor):
        if data.dim() == 1:
            data = data[None, :]
        lens = [data.shape[1]] * data.shape[0] if data_len is None else [int(x) for x in data_len]
        wavs = [data[i, :lens[i]] for i in range(data.shape[0])]
    elif isinstance(data, (list, tuple)):
        wavs = [torch.from_numpy(d) if isinstance(d, np.ndarray) else d for d in data]
        wavs = [w.reshape(-1) for w in wavs]
        lens = [int(w.numel()) for w in wavs]
    else:
        raise TypeError(f"unsupported data type {type(data)}")
    if not wavs:
        raise ValueError("empty batch")
    known = getattr(frontend, "known", None)  # batched.KnownFeatures: this utterance's features exist already
    if known is not None and len(wavs) == 1:
        hit = known(lens[0])
        if hit is not None:
            return hit[0].to(torch.float32), hit[1].to(torch.int32)
    wav, starts = pack_waveforms(wavs, frontend.device)
    n = np.asarray(lens, dtype=np.int64)
    _, T = frontend.lengths(n)
    feats, cu = frontend.forward_packed(wav, starts, n, out_tmax=int(T.max()))
    return feats.to(torch.float32), torch.from_numpy(np.diff(cu).astype(np.int32))
