"""Round-trip accuracy of the CUDA Griffin-Lim transforms on a well-conditioned problem (consistent spectrogram + its true
phases: the iteration must reproduce the signal), next to a float32 numpy/pocketfft pipeline.  Diagnostic only."""
import json, os, sys
import numpy as np, scipy.fft
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import spectral_oracle as so
from unast_b200 import synth, utils as U

def snr(a, b):
    e = a.astype(np.float64) - b
    return float(10 * np.log10((a.astype(np.float64) ** 2).sum() / max((e ** 2).sum(), 1e-300)))

w = synth.speech_like(10.0, 80)
X = so.stft(w)                                   # [1025, T] complex64
S = np.abs(X).astype(np.float32)
ang = (X / np.maximum(S, 1e-30)).astype(np.complex64)
ang[S == 0] = 1.0
ref0 = so.istft(X)                               # float64 irfft inside
out = {}
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
def gl_ref(iters, f32):
    Xc = (S * ang).astype(np.complex64)
    for _ in range(iters):
        x = istft32(Xc) if f32 else so.istft(Xc)
        E = stft32(x) if f32 else so.stft(x)
        Xc = (S * (E / np.maximum(1e-8, np.abs(E)))).astype(np.complex64)
    return istft32(Xc) if f32 else so.istft(Xc)
import torch
for it in (0, 1, 4):
    r64 = gl_ref(it, False)
    r32 = gl_ref(it, True)
    g = U._gl_batch([np.ascontiguousarray(S.T)], it, 0.0, [np.ascontiguousarray(ang.T)], 0.0, False, False)[0]
    out[str(it)] = {"numpy_f32_vs_f64": snr(r64, r32), "cuda_vs_f64": snr(r64, g), "cuda_vs_signal": snr(w[:g.size], g)}
print(json.dumps(out, indent=1))
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "gl_numerics.json"), "w"), indent=1)

# ---- zero-phase start (ill-conditioned: |E| is tiny where S is not), one iteration ----
(mel, mag), = U.spectrograms_from_waves([w], do_trim=False)
S2 = so.denormalise(mag)                                     # [1025, T] float32
ang1 = np.ones_like(S2, dtype=np.complex64)
def gl_zero(S_, iters, f32):
    Xc = S_.astype(np.complex64)
    for _ in range(iters):
        x = istft32(Xc) if f32 else so.istft(Xc)
        E = stft32(x) if f32 else so.stft(x)
        Xc = (S_ * (E / np.maximum(1e-8, np.abs(E)))).astype(np.complex64)
    return (istft32(Xc) if f32 else so.istft(Xc)), E
wss = so.window_sumsquare(S2.shape[1])
zero = {}
for it in (1, 2):
    r64, E64 = gl_zero(S2, it, False)
    r32, E32 = gl_zero(S2, it, True)
    g_direct = U._gl_batch([np.ascontiguousarray(S2.T)], it, 0.0, None, 0.0, False, False)[0]
    g_power = U._gl_batch([mag], it, 0.0, None, 1.2, False, False)[0]
    zero[str(it)] = {"numpy_f32_vs_f64": snr(r64, r32), "cuda_directS_vs_f64": snr(r64, g_direct), "cuda_power_vs_f64": snr(r64, g_power),
                     "cuda_directS_vs_numpy_f32": snr(r32, g_direct)}
print(json.dumps(zero, indent=1))
out["zero_phase"] = zero
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "gl_numerics.json"), "w"), indent=1)
