"""Speed-optimised cut: where a W-fragment run (linked seams, one launch, one GPU) differs from the single-fragment run.  Diagnostic."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from unast_b200 import longform as LF, synth, utils as U
secs = int(sys.argv[1]) if len(sys.argv) > 1 else 600
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 100
w = np.concatenate([synth.speech_like(10.0, 80 + (3 * i) % 128) for i in range((secs + 9) // 10)])[:secs * 22050]
(mel, mag), = U.spectrograms_from_waves([w], do_trim=False)
T = mag.shape[0]
mag_d = torch.from_numpy(mag).cuda()
one = LF.griffinlim_fragments_emulated(mag_d, 1, iters).cpu().numpy().astype(np.float64)
for W in (2, 4, 8):
    y = LF.griffinlim_fragments_emulated(mag_d, W, iters).cpu().numpy().astype(np.float64)
    e = y - one
    print("W", W, "snr %.1f" % (10 * np.log10((one ** 2).sum() / (e ** 2).sum())), flush=True)
    blk = 275 * 50
    nb = len(e) // blk
    be = np.array([(e[i * blk:(i + 1) * blk] ** 2).sum() for i in range(nb)])
    top = np.argsort(-be)[:6]
    print("   error energy: total %.3e; top 50-frame blocks (frame index, share):" % be.sum(), [(int(i * 50), round(float(be[i] / be.sum()), 3)) for i in top],
          " rank boundaries at frames", [T * r // W for r in range(1, W)])
