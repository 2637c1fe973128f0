"""CPU oracle for the UNAST spectral signal path -- TEST INFRASTRUCTURE ONLY.

This file restates, in plain numpy/scipy, the arithmetic that the reference runs on
the host for ``utils.get_spectrograms`` / ``utils.spectrogram2wav`` /
``gl_vocoder.vocode``.  It is the checker for the CUDA path; it is never imported
by the product package (only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s cpu_baseline / ``--impl reference`` legs may use it).

PARITY UNPINNED: the reference holds no golden vector, known-answer test or
fixture for this path, and its numerics live in
un-vendored third-party code (librosa==0.7.1 pinned only by
src/Transformer-TTS/requirements.txt:3; numpy.fft; scipy.signal) that is not
installed in the build image.  What pins this restatement instead:
  * tests/golden/make_golden.py imports the reference's OWN ``src/utils.py`` and
    ``src/gl_vocoder.py`` orchestration code from /root/reference and runs it over a
    ``librosa`` stand-in that is built from torch.stft/torch.istft (float64) and
    torchaudio's Slaney filterbank -- i.e. an implementation that shares no code
    with this file; its outputs are committed under tests/golden/.
  * the two shape known-answers the reference does contain
    (notebooks/testing_prepostnet.ipynb:593 -> 809 frames = 222,200 samples;
    Transformer-TTS/samples/test.wav = 275*399 samples).

Reference lines followed (all paths relative to /root/reference):
  src/utils.py:235-278   get_spectrograms
  src/utils.py:281-306   spectrogram2wav
  src/utils.py:309-320   griffin_lim
  src/utils.py:323-328   invert_spectrogram
  src/gl_vocoder.py:10-34 vocode (librosa.griffinlim variant)
  src/audio_parameters.py:1-20 constants
librosa 0.7.1 semantics are written out in SURVEY.md Appendix A.
"""
from __future__ import annotations

import struct

import numpy as np
from scipy import signal as _signal

# ---- src/audio_parameters.py:1-20 -------------------------------------------------------
sr = 22050
n_fft = 2048
preemphasis = 0.97
frame_shift = 0.0125
frame_length = 0.05
hop_length = int(sr * frame_shift)   # 275
win_length = int(sr * frame_length)  # 1102
n_mels = 80
power = 1.2
max_db = 100
ref_db = 20
n_iter = 60

N_BINS = 1 + n_fft // 2


# ---- librosa.load (src/utils.py:245; Appendix A.1) --------------------------------------
def read_wav(path):
    """Minimal RIFF/WAVE reader: PCM16 / PCM32 / float32, mono or multi-channel.

    Mirrors soundfile->float32 as used by librosa.load: PCM16 -> x/32768, channels averaged
    (librosa.to_mono).  Returns (float32 mono, sample_rate)."""
    with open(path, "rb") as fh:
        b = fh.read()
    if b[:4] != b"RIFF" or b[8:12] != b"WAVE":
        raise ValueError("not a RIFF/WAVE file: %s" % path)
    i, fmt, data = 12, None, None
    while i + 8 <= len(b):
        cid, sz = b[i:i + 4], struct.unpack("<I", b[i + 4:i + 8])[0]
        if cid == b"fmt ":
            fmt = struct.unpack("<HHIIHH", b[i + 8:i + 24])
        elif cid == b"data":
            data = b[i + 8:i + 8 + sz]
        i += 8 + sz + (sz & 1)
    if fmt is None or data is None:
        raise ValueError("missing fmt/data chunk: %s" % path)
    tag, nch, rate, _, _, bits = fmt
    if tag == 1 and bits == 16:
        y = np.frombuffer(data, "<i2").astype(np.float32) / np.float32(32768.0)
    elif tag == 1 and bits == 32:
        y = (np.frombuffer(data, "<i4").astype(np.float64) / 2147483648.0).astype(np.float32)
    elif tag == 3 and bits == 32:
        y = np.frombuffer(data, "<f4").astype(np.float32)
    else:
        raise ValueError("unsupported WAV encoding tag=%d bits=%d" % (tag, bits))
    if nch > 1:
        y = y.reshape(-1, nch).mean(axis=1).astype(np.float32)
    return y, rate


def load(path, sr_expected=sr):
    """librosa.load(path, sr=22050) for files already at 22,050 Hz (no resampling path)."""
    y, rate = read_wav(path)
    if rate != sr_expected:
        raise ValueError("resampling is out of scope: file sr=%d, expected %d" % (rate, sr_expected))
    return y, rate


# ---- windows / framing helpers (Appendix A.2) -------------------------------------------
def hann_periodic(n=win_length):
    """scipy.signal.get_window('hann', n, fftbins=True), float64."""
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / n)


def pad_center(w, size):
    lpad = (size - len(w)) // 2
    return np.pad(w, (lpad, size - len(w) - lpad), mode="constant")


def _frame(y, frame_len, hop):
    n_frames = 1 + (len(y) - frame_len) // hop
    idx = np.arange(frame_len)[:, None] + hop * np.arange(n_frames)[None, :]
    return y[idx]


# ---- librosa.effects.trim (src/utils.py:248,304; Appendix A.5) --------------------------
def trim_bounds(y, top_db=60, frame_len=2048, hop=512):
    """Returns (start, end) exactly as librosa.effects.trim would slice y."""
    y = np.asarray(y)
    ypad = np.pad(y, frame_len // 2, mode="reflect")
    x = _frame(ypad, frame_len, hop)
    powr = np.mean(np.abs(x) ** 2, axis=0, keepdims=True)
    mse = np.sqrt(powr) ** 2
    mse = mse.squeeze()
    log_spec = 10.0 * np.log10(np.maximum(1e-10, mse))
    log_spec = log_spec - 10.0 * np.log10(np.maximum(1e-10, np.max(mse)))
    nz = np.flatnonzero(log_spec > -top_db)
    if nz.size > 0:
        start = int(nz[0]) * hop
        end = min(y.shape[-1], (int(nz[-1]) + 1) * hop)
    else:
        start, end = 0, 0
    return start, end


def trim(y, top_db=60):
    s, e = trim_bounds(y, top_db)
    return y[s:e], np.asarray([s, e])


# ---- librosa.stft / istft (src/utils.py:254-257,314,328; Appendix A.2/A.3) ---------------
def stft(y, nfft=n_fft, hop=hop_length, win=win_length):
    w = pad_center(hann_periodic(win), nfft).reshape(-1, 1)       # float64
    ypad = np.pad(np.asarray(y), nfft // 2, mode="reflect")
    frames = _frame(ypad, nfft, hop)                               # (nfft, T), dtype of y
    out = np.empty((1 + nfft // 2, frames.shape[1]), dtype=np.complex64, order="F")
    # numpy >= 2.0 transforms float32 / complex64 input in single precision; the reference's numpy 1.19 (requirements.txt:2)
    # always transforms in float64 -- force it (w is float64, so the product already is)
    out[:] = np.fft.rfft((w * frames).astype(np.float64, copy=False), axis=0)   # float64 -> complex64
    return out


def window_sumsquare(n_frames, hop=hop_length, win=win_length, nfft=n_fft, dtype=np.float32):
    n = nfft + hop * (n_frames - 1)
    x = np.zeros(n, dtype=dtype)
    win_sq = pad_center(hann_periodic(win) ** 2, nfft)
    for i in range(n_frames):
        s = i * hop
        x[s:min(n, s + nfft)] += win_sq[:max(0, min(nfft, n - s))]
    return x


def istft(X, hop=hop_length, win=win_length, dtype=np.float32):
    nfft = 2 * (X.shape[0] - 1)
    T = X.shape[1]
    w = pad_center(hann_periodic(win), nfft)[:, None]
    y = np.zeros(nfft + hop * (T - 1), dtype=dtype)
    # float64 irfft as under the reference's numpy 1.19 (numpy >= 2.0 would run a complex64 input in single precision)
    ytmp = w * np.fft.irfft(np.asarray(X).astype(np.complex128), axis=0)
    for t in range(T):
        s = t * hop
        y[s:s + nfft] += ytmp[:, t]
    wss = window_sumsquare(T, hop, win, nfft, dtype=dtype)
    nzi = wss > np.finfo(dtype).tiny
    y[nzi] /= wss[nzi]
    return y[nfft // 2:-(nfft // 2)]


# ---- librosa.filters.mel (src/utils.py:263; Appendix A.4) --------------------------------
def _hz_to_mel(f):
    f = np.asanyarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    if f.ndim:
        lt = f >= min_log_hz
        mels[lt] = min_log_mel + np.log(f[lt] / min_log_hz) / logstep
    elif f >= min_log_hz:
        mels = min_log_mel + np.log(f / min_log_hz) / logstep
    return mels


def _mel_to_hz(m):
    m = np.asanyarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    freqs = f_sp * m
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    lt = m >= min_log_mel
    freqs[lt] = min_log_hz * np.exp(logstep * (m[lt] - min_log_mel))
    return freqs


def mel_filterbank(rate=sr, nfft=n_fft, nmels=n_mels):
    fmax = float(rate) / 2
    weights = np.zeros((nmels, 1 + nfft // 2), dtype=np.float32)
    fftfreqs = np.linspace(0, float(rate) / 2, 1 + nfft // 2, endpoint=True)
    mel_f = _mel_to_hz(np.linspace(_hz_to_mel(0.0), _hz_to_mel(fmax), nmels + 2))
    fdiff = np.diff(mel_f)
    ramps = np.subtract.outer(mel_f, fftfreqs)
    for i in range(nmels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:nmels + 2] - mel_f[:nmels])
    weights *= enorm[:, np.newaxis]
    return weights


_MEL_BASIS = None


def _mel_basis():
    # The reference rebuilds this on every call (src/utils.py:263); the values are constant.
    global _MEL_BASIS
    if _MEL_BASIS is None:
        _MEL_BASIS = mel_filterbank()
    return _MEL_BASIS


# ---- forward path (src/utils.py:235-278) -------------------------------------------------
def spectrograms_from_wave(y, do_trim=True, return_bounds=False):
    """get_spectrograms after the file has been decoded to float32 mono."""
    y = np.asarray(y, dtype=np.float32)
    bounds = (0, len(y))
    if do_trim:
        y, b = trim(y)                                             # utils.py:248
        bounds = (int(b[0]), int(b[1]))
    y = np.append(y[0], y[1:] - np.float32(preemphasis) * y[:-1])  # utils.py:251 (float32)
    linear = stft(y)                                               # utils.py:254-257
    mag = np.abs(linear)                                           # utils.py:260
    mel = np.dot(_mel_basis(), mag)                                # utils.py:263-264
    mel = 20 * np.log10(np.maximum(1e-5, mel))                     # utils.py:267
    mag = 20 * np.log10(np.maximum(1e-5, mag))                     # utils.py:268
    mel = np.clip((mel - ref_db + max_db) / max_db, 1e-8, 1)       # utils.py:271
    mag = np.clip((mag - ref_db + max_db) / max_db, 1e-8, 1)       # utils.py:272
    mel = mel.T.astype(np.float32)                                 # utils.py:275
    mag = mag.T.astype(np.float32)                                 # utils.py:276
    if return_bounds:
        return mel, mag, bounds
    return mel, mag


def get_spectrograms(fpath):
    y, _ = load(fpath)                                             # utils.py:245
    return spectrograms_from_wave(y)


# ---- inverse path (src/utils.py:281-328, src/gl_vocoder.py:17-32) ------------------------
def denormalise(mag):
    """utils.py:289-298: [T,1025] normalised -> S[1025,T] = (10^((clip*100-100+20)*0.05))^1.2."""
    mag = np.asarray(mag, dtype=np.float32).T
    # float32 throughout, exactly the reference's statements (utils.py:292-298).  NOTE: the zero-phase start of
    # utils.griffin_lim is ill-conditioned -- one ulp of S moves the 60-iteration waveform by about -72 dB (measured by
    # swapping numpy's powf for a correctly rounded one) -- so the last bit of these powf calls, which
    # differs between numpy versions / libms, is visible in Griffin-Lim's output at that level.
    mag = (np.clip(mag, 0, 1) * max_db) - max_db + ref_db
    mag = np.power(10.0, mag * 0.05)
    return (mag ** power).astype(np.float32)


def invert_spectrogram(X):
    return istft(X, hop_length, win_length)                        # utils.py:328


def griffin_lim(S, iters=None, angles0=None):
    """utils.py:309-320 (classic GL, zero-phase start).  angles0 (complex, |.|=1) optionally
    replaces the zero-phase start so both GL variants can share seeded initial phases."""
    iters = n_iter if iters is None else iters
    X_best = S.astype(np.complex64) * angles0.astype(np.complex64) if angles0 is not None else S.copy()
    for _ in range(iters):
        X_t = invert_spectrogram(X_best)
        est = stft(X_t)
        phase = est / np.maximum(1e-8, np.abs(est))
        X_best = S * phase
    X_t = invert_spectrogram(X_best)
    return np.real(X_t)


def librosa_griffinlim(S, iters=32, momentum=0.99, angles0=None, random_state=None):
    """librosa.griffinlim as called at gl_vocoder.py:26 (Appendix A.6)."""
    if angles0 is None:
        rng = np.random if random_state is None else np.random.RandomState(seed=random_state)
        angles0 = np.exp(2j * np.pi * rng.rand(*S.shape))
    angles = np.empty(S.shape, dtype=np.complex64)
    angles[:] = angles0
    rebuilt = 0.0
    for _ in range(iters):
        tprev = rebuilt
        inverse = istft(S * angles)
        rebuilt = stft(inverse)
        angles[:] = rebuilt - (momentum / (1 + momentum)) * tprev
        angles[:] /= np.abs(angles) + 1e-16
    return istft(S * angles)


def deemphasis(wav):
    return _signal.lfilter([1], [1, -preemphasis], wav)            # utils.py:301 (float64)


def spectrogram2wav(mag, iters=None, angles0=None, return_untrimmed=False):
    S = denormalise(mag)
    wav = griffin_lim(S, iters, angles0)                           # utils.py:298
    wav = deemphasis(wav)                                          # utils.py:301
    if return_untrimmed:
        return wav.astype(np.float32)
    wav, _ = trim(wav)                                             # utils.py:304
    return wav.astype(np.float32)                                  # utils.py:306


def vocode_one(mag, iters=32, momentum=0.99, angles0=None, random_state=None, return_untrimmed=False):
    """Body of the loop at gl_vocoder.py:14-32 for one utterance."""
    S = denormalise(mag)
    wav = librosa_griffinlim(S, iters, momentum, angles0, random_state)
    wav = deemphasis(wav)
    if return_untrimmed:
        return wav.astype(np.float32)
    wav, _ = trim(wav)
    return wav


def spectral_convergence(y, S):
    """|| |STFT(y)| - S ||_F / ||S||_F  (SURVEY.md 8(d) parity gate)."""
    E = np.abs(stft(np.asarray(y, dtype=np.float32)))
    T = min(E.shape[1], S.shape[1])
    return float(np.linalg.norm(E[:, :T] - S[:, :T]) / np.linalg.norm(S[:, :T]))


# ---- loader side: src/preprocess.py:82-121, :149-155, :176-182 restated in numpy -------------
def _pad_mel(inputs):
    """src/preprocess.py:176-182"""
    max_len = max((x.shape[0] for x in inputs))
    return np.stack([np.pad(x, [[0, max_len - x.shape[0]], [0, 0]], mode="constant", constant_values=0) for x in inputs])


def _prepare_data(inputs):
    """src/preprocess.py:149-155"""
    max_len = max((len(x) for x in inputs))
    return np.stack([np.pad(x, (0, max_len - x.shape[0]), mode="constant", constant_values=0) for x in inputs])


def collate_fn_transformer(batch):
    """src/preprocess.py:82-121 line by line on items {'text', 'mel', 'text_length', 'mel_length'[, 'fname']};
    returns numpy arrays (text int64, mel float32, text_length int64, mel_length int64)[, fnames]."""
    text = [d["text"] for d in batch]
    mel = [d["mel"] for d in batch]
    mel_length = [d["mel_length"] for d in batch]
    text_length = [d["text_length"] for d in batch]
    fnames = None
    if "fname" in batch[0]:
        fnames = [d["fname"] for d in batch]
        fnames = [i for i, _ in sorted(zip(fnames, text_length), key=lambda x: x[1], reverse=True)]
    text = [i for i, _ in sorted(zip(text, text_length), key=lambda x: x[1], reverse=True)]
    mel = [i for i, _ in sorted(zip(mel, text_length), key=lambda x: x[1], reverse=True)]
    mel_length = [i for i, _ in sorted(zip(mel_length, text_length), key=lambda x: x[1], reverse=True)]
    text_length = sorted(text_length, reverse=True)
    text = _prepare_data(text).astype(np.int32)
    mel = _pad_mel(mel)
    out = (text.astype(np.int64), mel.astype(np.float32), np.asarray(text_length, np.int64), np.asarray(mel_length, np.int64))
    return (out, fnames) if fnames is not None else out


def loader_items_from_waves(texts, waves, fnames=None):
    """LJDatasets.__getitem__ (src/preprocess.py:34-51) with the mel that prepare_data.py would have written for `wave`."""
    items = []
    for i, (tx, w) in enumerate(zip(texts, waves)):
        mel, _ = spectrograms_from_wave(w)
        d = {"text": np.asarray(tx, np.int32), "mel": mel, "text_length": len(tx), "mel_length": mel.shape[0]}
        if fnames is not None:
            d["fname"] = fnames[i]
        items.append(d)
    return items


# ---- next row: src/utils.py:40-75, specaugment / noise_fn ------------------------------------
# PINNED: tests/golden/make_augment_golden.py runs the reference's own two functions (torch only) and commits their
# outputs as tests/golden/augment.npz; tests/test_oracle_golden.py checks this restatement against them bit for bit.
def specaugment_spans(mel_len, freq_mask=20, time_mask=100):
    """The random draws of utils.specaugment (utils.py:55-68), in its order: torch.randint for the span lengths of the
    whole batch, then random.randrange per utterance.  Returns int32 [B, 4] rows (f_zero, f, t_zero, t)."""
    import random
    import torch
    mel_len = torch.as_tensor(np.asarray(mel_len))
    freq_len = torch.randint(0, freq_mask, mel_len.shape)                   # utils.py:55
    time_len = torch.randint(0, time_mask, mel_len.shape)                   # utils.py:56
    spans = np.zeros((mel_len.shape[0], 4), np.int32)
    for i in range(freq_len.shape[0]):
        n = int(mel_len[i])
        f, t = int(freq_len[i]), int(time_len[i])
        if n - t <= 0:                                                      # utils.py:63-65
            t = random.randrange(0, n // 2)
        f_zero = random.randrange(0, n - f)                                 # utils.py:66
        t_zero = random.randrange(0, n - t)                                 # utils.py:67
        spans[i] = (f_zero, f, t_zero, t)
    return spans


def specaugment_apply(mel, spans, replace_with_zero=False):
    """utils.py:58-73 given the draws: mean of the padded slab first, then two ROW spans (``res[i][:][a:b]`` slices the
    time axis) filled with it (or 0).  float32 throughout; the mean is numpy's pairwise float32 sum / count, which
    torch's float32 mean matches to rounding (the golden test allows 1 ulp of the fill value)."""
    res = np.array(mel, dtype=np.float32, copy=True)
    for i, (f_zero, f, t_zero, t) in enumerate(np.asarray(spans)):
        fill = np.float32(0.0) if replace_with_zero else res[i].mean(dtype=np.float32)
        res[i, f_zero:f_zero + f, :] = fill
        res[i, t_zero:t_zero + t, :] = fill
    return res


def noise_apply(to_noise, zero_mask):
    """utils.py:46-49 given the Bernoulli draw: ``to_noise * zero_mask.unsqueeze(-1)``."""
    return np.asarray(to_noise, np.float32) * np.asarray(zero_mask, np.float32)[..., None]
