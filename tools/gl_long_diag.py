"""Per-second SNR of the CUDA Griffin-Lim against the oracle on a 60 s clip (diagnostic; writes gpurun_out/gl_long_diag.json)."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import spectral_oracle as so
from unast_b200 import synth, utils as U

def snr(a, b):
    e = a.astype(np.float64) - b
    return float(10 * np.log10((a.astype(np.float64) ** 2).sum() / max((e ** 2).sum(), 1e-300)))

iters = int(sys.argv[1]) if len(sys.argv) > 1 else 30
w = np.concatenate([synth.speech_like(10.0, 80 + i) for i in range(6)])
(mel, mag), = U.spectrograms_from_waves([w], do_trim=False)
out = {}
for it in (1, 5, iters):
    y, = U.spectrogram2wav_batch([mag], n_iter=it, do_trim=False)
    ref = so.spectrogram2wav(mag, iters=it, return_untrimmed=True)
    n = ref.size // 60
    out[str(it)] = {"total": snr(ref, y), "per_s": [round(snr(ref[i * n:(i + 1) * n], y[i * n:(i + 1) * n]), 1) for i in range(60)]}
    # the same clip cut into 6 utterances of 10 s (no long utterance): is it the length or the content?
mags6 = [m for _, m in U.spectrograms_from_waves([synth.speech_like(10.0, 80 + i) for i in range(6)], do_trim=False)]
ys = U.spectrogram2wav_batch(mags6, n_iter=iters, do_trim=False)
out["six_short"] = [snr(so.spectrogram2wav(m, iters=iters, return_untrimmed=True), y) for m, y in zip(mags6, ys)]
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "gl_long_diag.json"), "w"), indent=1)
print(json.dumps(out))
