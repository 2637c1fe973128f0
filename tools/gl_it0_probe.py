import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import spectral_oracle as so
from unast_b200 import synth, utils as U
w = synth.speech_like(10.0, 80)
mel, mag = so.spectrograms_from_wave(w, do_trim=False)
S = so.denormalise(mag)
r64 = so.istft(S.astype(np.complex64)).astype(np.float64)
g = U._gl_batch([np.ascontiguousarray(S.T)], 0, 0.0, None, 0.0, False, False)[0].astype(np.float64)
e = r64 - g
m = (e.size // 275) * 275
E = e[:m].reshape(-1, 275)[8:-8]
R = r64[:m].reshape(-1, 275)[8:-8]
prof = (E ** 2).mean(0)
print("total snr", 10 * np.log10((r64 ** 2).sum() / (e ** 2).sum()))
print("top hop phases", np.argsort(-prof)[:12].tolist(), (np.sort(prof)[::-1][:12] / prof.mean()).round(1).tolist())
print("signal top phases", np.argsort(-(R ** 2).mean(0))[:6].tolist())
# is the error proportional to the signal (a gain error)?  least squares gain per hop phase
gain = (E * R).sum(0) / (R * R).sum(0)
print("relative gain error by hop phase: max |.|", np.abs(gain).max(), "at", int(np.abs(gain).argmax()), "median", float(np.median(np.abs(gain))))
res = E - gain[None, :] * R
print("snr after removing per-phase gain", 10 * np.log10((R ** 2).sum() / (res ** 2).sum()))
np.save(os.path.join(ROOT, "gpurun_out", "it0_gain.npy"), gain)
