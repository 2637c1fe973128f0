"""Error of the float32 front-end mode against the float64 parity mode (itself within 2e-7 of the oracle) by how far a bin
sits below its frame's peak -- the measurement that closes the "can a float32 FFT meet the 1e-4 gate" question.
Run on a B200: python tools/fe_f32_histogram.py [n_utts] > profiles/r3_fe_f32_error_histogram.txt"""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from unast_b200 import synth
from unast_b200.engine import get_engine

n_utts = int(sys.argv[1]) if len(sys.argv) > 1 else 512
eng = get_engine()
durs = synth.ljspeech_durations(13100, seed=1)[:: max(1, 13100 // n_utts)][:n_utts]
waves = [synth.to_pcm16(synth.speech_like(float(d), 1000 + i)) for i, d in enumerate(durs)]
g = np.load(os.path.join(ROOT, "tests", "golden", "real_test_wav.npz"))
sets = {"synthetic C2 sample (%d utterances, distinct seeds)" % len(waves): waves, "reference samples/test.wav (real speech, float32 input)": [g["wave"]]}
edges = [0, -20, -40, -60, -80, -90, -100, -110, -120, -140, -200]
for name, ws in sets.items():
    ws = [np.ascontiguousarray(w) for w in ws]
    lens = np.array([w.size for w in ws]); off = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)])).cuda()
    wave = torch.from_numpy(np.concatenate(ws)).cuda()
    bounds = eng.trim_bounds(wave, off, int(lens.max())); fo = eng.frame_offsets(bounds); eng.check()
    F = int(fo.cpu()[-1])
    _, m64 = eng.frontend(wave, off, bounds, fo, F, precision="f64")
    _, m32 = eng.frontend(wave, off, bounds, fo, F, precision="f32")
    d = (m32 - m64).abs()
    db = m64 * 100.0 - 80.0                                   # 20 log10 |X| (clipped at -80)
    lvl = db - db.max(dim=1, keepdim=True).values              # dB below the frame's peak
    live = m64 > 1e-8                                          # bins above the clip floor
    print("# %s: %d frames, %d bins above the clip floor, max |d| %.3e, bins > 1e-4: %d, > 5e-5: %d" % (
        name, F, int(live.sum()), float(d.max()), int((d > 1e-4).sum()), int((d > 5e-5).sum())))
    print("# dB below frame peak      bins        max |d|     mean |d|    > 5e-5     > 1e-4")
    for hi, lo in zip(edges[:-1], edges[1:]):
        m = live & (lvl <= hi) & (lvl > lo)
        n = int(m.sum())
        if n:
            dm = d[m]
            print("  (%5d, %5d]   %12d   %10.3e  %10.3e  %8d  %8d" % (lo, hi, n, float(dm.max()), float(dm.mean()), int((dm > 5e-5).sum()), int((dm > 1e-4).sum())))
    print()
