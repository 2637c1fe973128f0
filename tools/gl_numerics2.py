"""Where does the CUDA Griffin-Lim lose accuracy on the zero-phase start?  Error profile over time for it = 0, 1."""
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
mel, mag = so.spectrograms_from_wave(w, do_trim=False)
S = so.denormalise(mag)
out = {}
for it in (0, 1, 2):
    X = S.astype(np.complex64)
    for _ in range(it):
        E = so.stft(so.istft(X)); X = (S * (E / np.maximum(1e-8, np.abs(E)))).astype(np.complex64)
    r64 = so.istft(X)
    g = U._gl_batch([np.ascontiguousarray(S.T)], it, 0.0, None, 0.0, False, False)[0]
    n = r64.size
    blocks = [(0, 1100), (1100, 2200), (2200, 4400), (4400, n - 4400), (n - 4400, n - 2200), (n - 2200, n - 1100), (n - 1100, n)]
    err = (r64.astype(np.float64) - g) ** 2
    sig = r64.astype(np.float64) ** 2
    out[str(it)] = {"total": snr(r64, g), "blocks": [[a, b, round(10 * np.log10(sig[a:b].sum() / max(err[a:b].sum(), 1e-300)), 1),
                                                       float(err[a:b].mean()), float(sig[a:b].mean())] for a, b in blocks]}
    # error energy per 275-sample hop modulo position (periodic structure -> table problem)
    m = (n // 275) * 275
    e2 = err[:m].reshape(-1, 275)[8:-8]
    out[str(it)]["err_by_hop_phase_max_over_mean"] = float(e2.mean(0).max() / e2.mean())
    out[str(it)]["err_frames_top"] = [int(i) for i in np.argsort(-e2.mean(1))[:8]]
print(json.dumps({k: {"total": v["total"], "hop": v["err_by_hop_phase_max_over_mean"], "blocks": [b[2] for b in v["blocks"]]} for k, v in out.items()}))
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "gl_numerics2.json"), "w"), indent=1)
