"""Accuracy of the product FFT cores by bin magnitude (debug entry points), next to pocketfft float32.  Diagnostic."""
import json, os, sys
import numpy as np, scipy.fft
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from oracle import spectral_oracle as so
from unast_b200 import synth, utils as U
eng = U.get_engine()
w = synth.speech_like(10.0, 80)
mel, mag = so.spectrograms_from_wave(w, do_trim=False)
S = so.denormalise(mag).T.astype(np.float32)
fr64 = np.roll(np.fft.irfft(S.astype(np.float64), n=2048, axis=1), -473, axis=1)[:, :1102]
win = so.hann_periodic()
xs = np.zeros((S.shape[0], 2048), np.float32)
xs[:, :1102] = (fr64 * win * win / 1.5).astype(np.float32)
E64 = np.fft.rfft(xs.astype(np.float64), axis=1)
E_p = scipy.fft.rfft(xs, axis=1)
E_g = eng.debug_rfft(torch.from_numpy(xs).cuda(), "f32").cpu().numpy()
E_d = eng.debug_rfft(torch.from_numpy(xs).cuda(), "f64").cpu().numpy()
aw = np.abs(E64)
out = {"fwd": []}
for lo, hi in ((0, 1e-6), (1e-6, 1e-5), (1e-5, 1e-4), (1e-4, 1e-3), (1e-3, 1)):
    m = (aw >= lo) & (aw < hi)
    if m.sum():
        out["fwd"].append([lo, hi, int(m.sum()), float(np.abs(E_g - E64)[m].mean() / aw[m].mean()), float(np.abs(E_p - E64)[m].mean() / aw[m].mean()),
                           float(np.abs(E_d - E64)[m].mean() / aw[m].mean())])
# inverse on the zero-phase spectrum (shifted): taps region
k = np.arange(1025)
phi = np.exp(2j * np.pi * ((473 * k) % 2048) / 2048)
X0 = (S * phi).astype(np.complex64)
ref = np.fft.irfft(X0.astype(np.complex128), n=2048, axis=1)[:, :1102]
g = eng.debug_irfft(torch.from_numpy(X0).cuda()).cpu().numpy()
p = scipy.fft.irfft(X0, n=2048, axis=1)[:, :1102]
def snr(a, b):
    e = a.astype(np.float64) - b
    return float(10 * np.log10((a.astype(np.float64) ** 2).sum() / max((e ** 2).sum(), 1e-300)))
out["inv_zero_phase_taps_snr"] = {"cuda": snr(ref, g), "pocketfft32": snr(ref, p)}
print(json.dumps(out, indent=1))
