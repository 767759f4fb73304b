"""TEST INFRASTRUCTURE (oracle) - CPU restatement of the reference's audio ingest step, SURVEY.md §8(f) rank 2:
``torchaudio.transforms.Resample(sr, 16000)(audio.float())`` (/root/reference/infer.py:320-330, infer_kaldidata.py:454-455).
The arithmetic lives in torchaudio (TA/functional/functional.py:1305-1432): a polyphase windowed-sinc filter bank
(``sinc_interp_hann``, lowpass_filter_width 6, rolloff 0.99) applied as a strided conv1d.  Pinned by
tests/golden/resample_golden.npz (outputs of the real torchaudio 2.11 transform, generator tests/golden/make_golden_resample.py).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module."""
import math

import numpy as np


def sinc_resample_kernel(orig_freq: int, new_freq: int, lowpass_filter_width: int = 6, rolloff: float = 0.99):
    """TA/functional/functional.py:1305-1402 with dtype=None (what transforms.Resample passes): the table is built in
    float64 - except the phase offsets j/new, which torch computes as an int64 tensor divided by an int, i.e. in float32 -
    and rounded to float32 at the end.  Returns (kernels [new, 2*width+orig] float32, width, orig, new) with the
    frequencies reduced by their gcd."""
    g = math.gcd(int(orig_freq), int(new_freq))
    orig, new = int(orig_freq) // g, int(new_freq) // g
    base = min(orig, new) * rolloff
    width = math.ceil(lowpass_filter_width * orig / base)
    idx = np.arange(-width, width + orig, dtype=np.float64)[None, :] / orig
    phase = (np.arange(0, -new, -1, dtype=np.int64).astype(np.float32) / np.float32(new)).astype(np.float64)[:, None]
    t = (phase + idx) * base
    t = np.clip(t, -lowpass_filter_width, lowpass_filter_width)
    window = np.cos(t * math.pi / lowpass_filter_width / 2) ** 2
    t = t * math.pi
    scale = base / orig
    with np.errstate(invalid="ignore", divide="ignore"):
        k = np.where(t == 0, 1.0, np.sin(t) / t)
    k = k * (window * scale)
    return k.astype(np.float32), width, orig, new


def resample(wave: np.ndarray, orig_freq: int, new_freq: int) -> np.ndarray:
    """TA/functional/functional.py:1405-1432: pad (width, width+orig), conv1d with stride orig, interleave the `new`
    phases, crop to ceil(new*L/orig).  fp32 like the reference (accumulation order is conv1d's, so compare with a
    tolerance, not bit-for-bit)."""
    x = np.asarray(wave, dtype=np.float32).reshape(-1)
    if int(orig_freq) == int(new_freq):
        return x.copy()
    k, width, orig, new = sinc_resample_kernel(orig_freq, new_freq)
    L = x.shape[0]
    xp = np.concatenate([np.zeros(width, np.float32), x, np.zeros(width + orig, np.float32)])
    K = k.shape[1]
    n_frames = (xp.shape[0] - K) // orig + 1
    frames = np.lib.stride_tricks.as_strided(xp, shape=(n_frames, K), strides=(xp.strides[0] * orig, xp.strides[0]))
    out = (frames.astype(np.float64) @ k.astype(np.float64).T).astype(np.float32)  # [frames, new]
    target = -(-new * L // orig)
    return out.reshape(-1)[:target]


def pcm_to_mono_f32(wave: np.ndarray, scale: float = 1.0) -> np.ndarray:
    """Raw PCM ``[L]`` or ``[L, C]`` (frames x channels, int16 / float32 / float64) -> mono float32.

    * ``torch.from_numpy(audio).float()`` (/root/reference/infer.py:320-324): value-preserving cast to float32 (int16 exact,
      float64 rounded to nearest); scale 1.  Pinned in tests/test_ingest.py against torch itself.
    * ``librosa.load(path, sr=16000)`` (/root/reference/infer_kaldidata.py:101): soundfile hands 16-bit files over as
      float32 ``x / 32768`` (scale 1/32768) and ``librosa.to_mono`` is ``np.mean`` over the channel axis, i.e. the channels
      summed in order in float32 and divided by their count.  librosa / soundfile are not on this box: this part restates
      their documented arithmetic (librosa 0.10 ``core/audio.py:to_mono``) and is parity-unpinned."""
    x = np.asarray(wave)
    if x.ndim == 1:
        x = x[:, None]
    v = x.astype(np.float32) * np.float32(scale)
    s = v[:, 0].copy()
    for c in range(1, v.shape[1]):
        s = s + v[:, c]
    return (s / np.float32(v.shape[1])).astype(np.float32) if v.shape[1] > 1 else s
