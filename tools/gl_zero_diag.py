"""Zero-phase Griffin-Lim start, iteration by iteration: CUDA and two numpy float32 pipelines (all float32 / initial inverse in
float64) against the float64 oracle.  Diagnostic only."""
import json, os, sys
import numpy as np, scipy.fft
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import spectral_oracle as so
from unast_b200 import synth, utils as U

def snr(a, b):
    e = a.astype(np.float64) - b
    return round(float(10 * np.log10((a.astype(np.float64) ** 2).sum() / max((e ** 2).sum(), 1e-300))), 1)

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1021
w = synth.speech_like(1.5, seed)
(mel, mag), = U.spectrograms_from_waves([w])
S = so.denormalise(mag)                                      # [1025, T] float32
win = so.pad_center(so.hann_periodic(), 2048).astype(np.float32)
wss = so.window_sumsquare(S.shape[1])
def stft32(y):
    yp = np.pad(y.astype(np.float32), 1024, mode='reflect'); T = 1 + y.size // 275
    idx = np.arange(2048)[None, :] + 275 * np.arange(T)[:, None]
    return scipy.fft.rfft(yp[idx] * win[None, :], axis=1).T.astype(np.complex64)
def istft32(Xc):
    T = Xc.shape[1]
    fr = scipy.fft.irfft(Xc.T.astype(np.complex64), n=2048, axis=1).astype(np.float32) * win[None, :]
    y = np.zeros(2048 + 275 * (T - 1), np.float32)
    for t in range(T): y[275 * t:275 * t + 2048] += fr[t]
    nz = wss > np.finfo(np.float32).tiny; y[nz] /= wss[nz]
    return y[1024:-1024]
def run(iters, mode):                                        # 0: float64 oracle, 1: float32, 2: float32 after a float64 initial inverse
    Xc = S.astype(np.complex64)
    for i in range(iters):
        x = so.istft(Xc) if (mode == 0 or (mode == 2 and i == 0)) else istft32(Xc)
        E = so.stft(x) if mode == 0 else stft32(x)
        Xc = (S * (E / np.maximum(1e-8, np.abs(E)))).astype(np.complex64)
    return so.istft(Xc) if (mode == 0 or (mode == 2 and iters == 0)) else istft32(Xc)
out = {}
for it in (0, 1, 2, 5, 30, 100):
    r64 = run(it, 0)
    g = U._gl_batch([mag], it, 0.0, None, 1.2, False, False)[0]                        # S = gl_prepare_kernel(mag): correctly rounded powers
    gs = U._gl_batch([np.ascontiguousarray(S.T)], it, 0.0, None, 0.0, False, False)[0]  # the oracle's S (numpy's float32 power) passed through
    out[it] = {"cuda": snr(r64, g), "cuda_same_S": snr(r64, gs), "numpy_f32": snr(r64, run(it, 1)), "numpy_f32_init64": snr(r64, run(it, 2))}
    print(it, out[it], flush=True)
