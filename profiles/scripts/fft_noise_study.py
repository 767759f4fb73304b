"""Round-off study for the frontend FFT (CPU only, numpy): emulates the kernel's 512-pt real FFT (256-pt complex as two
radix-16 passes + split) step by step in fp32 and compares its error against fp64 truth with numpy's fp32 pocketfft
(= what torchaudio reaches).  Reports the error of the low bins, where pre-emphasis leaves ~1e-3 of the energy and
fp32 round-off from the loud bins is visible in the log-mel (tests/test_gpu_frontend.py bound)."""
import sys

import numpy as np

f32 = np.float32
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
NF = 3000


def window(n=400):
    k = np.arange(n)
    return ((0.5 - 0.5 * np.cos(2 * np.pi * k / (n - 1))).astype(f32).astype(np.float64) ** 0.85)


def frames64():
    x = (rng.random((NF, 400)) - 0.5) * 32768.0
    x = x.astype(f32).astype(np.float64)
    x = x - x.mean(axis=1, keepdims=True)
    prev = np.concatenate([x[:, :1], x[:, :-1]], axis=1)
    y = (x - 0.97 * prev) * window()[None]
    out = np.zeros((NF, 512))
    out[:, :400] = y
    return out


def cmul(a, b):
    # complex64 multiply with fp32 rounding of every op (no FMA): an upper bound of the kernel's error
    ar, ai, br, bi = a.real.astype(f32), a.imag.astype(f32), b.real.astype(f32), b.imag.astype(f32)
    return ((ar * br - ai * bi).astype(f32) + 1j * (ar * bi + ai * br).astype(f32)).astype(np.complex64)


def dft4(a, b, c, d):
    s0, s1, s2, s3 = a + c, a - c, b + d, b - d
    return s0 + s2, s1 - 1j * s3, s0 - s2, s1 + 1j * s3


def fft16(x):  # x: [..., 16] complex64 natural order -> X[q] natural order
    C1, S1, R = f32(0.92387953251128674), f32(0.38268343236508977), f32(0.70710678118654752)
    x = x.astype(np.complex64)
    A = np.empty(x.shape, np.complex64)
    for j0 in range(4):
        o = dft4(x[..., j0], x[..., 4 + j0], x[..., 8 + j0], x[..., 12 + j0])
        for q0 in range(4):
            A[..., 4 * q0 + j0] = o[q0]
    W = np.exp(-2j * np.pi * np.arange(16) / 16).astype(np.complex64)
    for q0 in range(1, 4):
        for j0 in range(1, 4):
            A[..., 4 * q0 + j0] = cmul(A[..., 4 * q0 + j0], np.broadcast_to(W[(q0 * j0) % 16], A[..., 0].shape))
    X = np.empty(x.shape, np.complex64)
    for q0 in range(4):
        o = dft4(A[..., 4 * q0 + 0], A[..., 4 * q0 + 1], A[..., 4 * q0 + 2], A[..., 4 * q0 + 3])
        for q1 in range(4):
            X[..., q0 + 4 * q1] = o[q1]
    return X


def kernel_rfft(fr):  # fr [NF,512] float32
    z = (fr[:, 0::2] + 1j * fr[:, 1::2]).astype(np.complex64)  # 256 complex points n
    # n = hl + 16 j  -> x[hl][j]
    x = z.reshape(NF, 16, 16).transpose(0, 2, 1)  # [NF, hl, j]
    X = fft16(x)  # [NF, hl, q]
    tw = np.exp(-2j * np.pi * (np.arange(16)[:, None] * np.arange(16)[None, :]) / 256).astype(np.complex64)
    X = cmul(X, np.broadcast_to(tw[None], X.shape))
    Y = fft16(X.transpose(0, 2, 1))  # [NF, q, p] over t=hl -> Z[q + 16 p]
    Z = Y.transpose(0, 2, 1).reshape(NF, 256)
    k = np.arange(0, 129)
    zk, zm = Z[:, k], np.conj(Z[:, (256 - k) % 256])
    xe = (f32(0.5) * (zk + zm)).astype(np.complex64)
    xo = (f32(0.5) * (-1j) * (zk - zm)).astype(np.complex64)
    w = np.exp(-2j * np.pi * k / 512).astype(np.complex64)
    t = cmul(xo, np.broadcast_to(w[None], xo.shape))
    lo = xe + t
    hi = np.conj(xe - t)  # X[256-k]
    out = np.empty((NF, 257), np.complex64)
    out[:, k] = lo
    out[:, 256 - k] = hi
    return out


fr64 = frames64()
truth = np.fft.rfft(fr64, axis=1)
fr32 = fr64.astype(f32)
truth32in = np.fft.rfft(fr32.astype(np.float64), axis=1)  # truth for the SAME fp32 frames
pocket = np.fft.rfft(fr32, axis=1)
assert pocket.dtype == np.complex64
mine = kernel_rfft(fr32)
for name, got in (("pocketfft f32", pocket), ("kernel emu f32", mine)):
    e = np.abs(got - truth32in)
    print(f"{name:16s} rms abs err: all bins {np.sqrt((e**2).mean()):.4f}  bins1-8 {np.sqrt((e[:, 1:9]**2).mean()):.4f}  "
          f"max {e.max():.3f};  rms |X| {np.sqrt((np.abs(truth32in)**2).mean()):.0f}  rms|X[1:3]| {np.sqrt((np.abs(truth32in[:,1:3])**2).mean()):.0f}")
    P, Pt = np.abs(got.astype(np.complex128)) ** 2, np.abs(truth32in) ** 2
    le = np.abs(np.log(P[:, 1:9]) - np.log(Pt[:, 1:9]))
    print(f"{'':16s} max |dlog P| bins1-8: {le.max():.2e}  (bin {1 + le.max(axis=0).argmax()})")
