"""Where the zero-iteration output (w * irfft(S), overlap-added, normalised) differs from the oracle.  Diagnostic only."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import spectral_oracle as so
from unast_b200 import synth, utils as U
w = synth.speech_like(1.5, 1021)
(mel, mag), = U.spectrograms_from_waves([w])
S = so.denormalise(mag)
ref = so.istft(S.astype(np.complex64)).astype(np.float64)
g = U._gl_batch([mag], 0, 0.0, None, 1.2, False, False)[0].astype(np.float64)
g2 = U._gl_batch([np.ascontiguousarray(S.T)], 0, 0.0, None, 0.0, False, False)[0].astype(np.float64)   # S passed through (power 0)
e = g - ref
print("T", S.shape[1], "len", ref.size, "snr all %.1f" % (10 * np.log10((ref ** 2).sum() / (e ** 2).sum())),
      "direct-S %.1f" % (10 * np.log10((ref ** 2).sum() / ((g2 - ref) ** 2).sum())))
nb = ref.size // 275
for b in list(range(0, 6)) + list(range(nb // 2, nb // 2 + 3)) + list(range(nb - 5, nb)):
    sl = slice(275 * b, 275 * (b + 1))
    print(b, "ref rms %.3e err rms %.3e  snr %.1f" % (np.sqrt((ref[sl] ** 2).mean()), np.sqrt((e[sl] ** 2).mean()),
          10 * np.log10((ref[sl] ** 2).sum() / max((e[sl] ** 2).sum(), 1e-300))))
i = np.argsort(-np.abs(e))[:8]
print("worst samples", [(int(k), float(ref[k]), float(g[k])) for k in i])
