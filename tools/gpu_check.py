"""Verbose GPU-vs-oracle diagnostics (prints error statistics; the pass/fail gates live in tests/).
Run on a B200: python tools/gpu_check.py [fft] [fe] [gl]"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import spectral_oracle as so  # noqa: E402
from unast_b200 import synth  # noqa: E402
from unast_b200 import utils as U  # noqa: E402
from unast_b200.engine import get_engine  # noqa: E402
from unast_b200.gl_vocoder import vocode_arrays  # noqa: E402

what = set(sys.argv[1:]) or {"fft", "fe", "gl"}
eng = get_engine()
print("engine up, SMs:", eng.num_sms, flush=True)


def stats(name, d):
    d = np.abs(np.asarray(d, dtype=np.float64)).ravel()
    print("  %-28s max %.3e  p99.99 %.3e  mean %.3e  n>1e-4 %d  n>5e-5 %d" %
          (name, d.max(), np.quantile(d, 0.9999), d.mean(), (d > 1e-4).sum(), (d > 5e-5).sum()), flush=True)


if "fft" in what:
    rng = np.random.RandomState(0)
    F = 37
    fr = np.zeros((F, 2048), np.float32)
    fr[:, :1102] = rng.randn(F, 1102).astype(np.float32)
    ref = np.fft.rfft(fr.astype(np.float64), axis=1)
    for prec in ("f32", "f64"):
        out = eng.debug_rfft(torch.from_numpy(fr).cuda(), prec).cpu().numpy()
        print("rfft %s: max abs err %.3e (ref max %.1f)" % (prec, np.abs(out - ref).max(), np.abs(ref).max()), flush=True)
    spec = ref.astype(np.complex64)
    back = eng.debug_irfft(torch.from_numpy(spec).cuda()).cpu().numpy()
    refb = np.fft.irfft(spec.astype(np.complex128), axis=1)[:, :1102]
    print("irfft f32: max abs err %.3e (ref max %.2f)" % (np.abs(back - refb).max(), np.abs(refb).max()), flush=True)

if "fe" in what:
    waves = [synth.speech_like(d, s) for d, s in [(10.0, 0), (3.3, 1), (1.0, 2), (6.57, 3)]]
    for prec in ("f64", "f32"):
        t0 = time.time()
        res = U.spectrograms_from_waves(waves, precision=prec, return_bounds=True)
        print("front end %s (%.2fs):" % (prec, time.time() - t0), flush=True)
        for w, (mel, mag, b) in zip(waves, res):
            omel, omag, ob = so.spectrograms_from_wave(w, return_bounds=True)
            print(" bounds gpu %s oracle %s shapes %s %s" % (b, ob, mel.shape, mag.shape))
            if mel.shape == omel.shape:
                stats("mag", mag - omag)
                stats("mel", mel - omel)
    # int16 input
    q = synth.to_pcm16(waves[1])
    mel, mag = U.spectrograms_from_waves([q], precision="f64")[0]
    omel, omag = so.spectrograms_from_wave(q.astype(np.float32) / np.float32(32768))
    print(" int16 input:")
    stats("mag", mag - omag)
    stats("mel", mel - omel)
    # padded loader layout + no-trim
    mb = eng.mel_basis()
    print(" mel basis vs oracle: max diff %.3e" % np.abs(mb - so.mel_filterbank()).max())

if "gl" in what:
    waves = [synth.speech_like(d, s) for d, s in [(2.0, 10), (1.2, 11), (3.1, 12)]]
    mags = [so.spectrograms_from_wave(w)[1] for w in waves]
    for iters in (0, 1, 5, 30):
        t0 = time.time()
        outs = U.spectrogram2wav_batch(mags, n_iter=iters, do_trim=False)
        dt = time.time() - t0
        for m, y in zip(mags, outs):
            ref = so.spectrogram2wav(m, iters=iters, return_untrimmed=True)
            err = y.astype(np.float64) - ref
            snr = 10 * np.log10((ref.astype(np.float64) ** 2).sum() / max((err ** 2).sum(), 1e-300))
            print(" classic it=%d T=%d len %d/%d SNR %.1f dB max|err| %.2e (ref max %.2f) %.2fs" %
                  (iters, m.shape[0], y.size, ref.size, snr, np.abs(err).max(), np.abs(ref).max(), dt), flush=True)
    # trim parity + spectral convergence
    outs = U.spectrogram2wav_batch(mags, n_iter=30)
    for m, y in zip(mags, outs):
        ref = so.spectrogram2wav(m, iters=30)
        S = so.denormalise(m)
        print(" trimmed len gpu %d oracle %d" % (y.size, ref.size))
    # momentum variant with seeded phases
    rs = np.random.RandomState(3)
    a0 = [np.exp(2j * np.pi * rs.rand(1025, m.shape[0])) for m in mags]
    outs = vocode_arrays(mags, n_iter=32, momentum=0.99, angles0=[a.T for a in a0], do_trim=False)
    for m, a, y in zip(mags, a0, outs):
        ref = so.vocode_one(m, iters=32, momentum=0.99, angles0=a, return_untrimmed=True)
        err = y.astype(np.float64) - ref
        snr = 10 * np.log10((ref.astype(np.float64) ** 2).sum() / max((err ** 2).sum(), 1e-300))
        print(" momentum it=32 T=%d SNR %.1f dB" % (m.shape[0], snr), flush=True)
print("done")
