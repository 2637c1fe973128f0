"""Drop-in for the spectral block of the reference's ``src/utils.py`` (lines 235-328).

Same names, signatures, return shapes / dtypes / memory order and error behaviour:

    get_spectrograms(fpath) -> (mel [T, 80] float32, mag [T, 1025] float32)        utils.py:235
    spectrogram2wav(mag)    -> wav [L] float32                                     utils.py:281
    griffin_lim(spectrogram [1025, T]) -> y [275 (T-1)]                            utils.py:309
    invert_spectrogram(spectrogram complex [1025, T]) -> y                         utils.py:323

Everything numeric runs in libunast_spectral.so on the GPU (no librosa, no numpy FFT, no fallback);
this file only moves bytes.  Batched variants (``*_batch``) exist because one utterance cannot fill a
B200 -- ``prepare_data`` and ``gl_vocoder`` use those.
"""
from __future__ import annotations

import numpy as np
import torch

from . import audio_parameters as ap
from .engine import HOP, NBINS, NMELS, get_engine
from .wavio import read_wav

# "f64" reproduces the reference's float64 FFT and meets the 1e-4 parity gate on every bin; "f32" is the
# fast mode (max error ~2.5e-4 on bins ~90 dB below the frame peak).  See DESIGN.md "precision".
DEFAULT_PRECISION = "f64"


def _as_wave_array(y):
    y = np.asarray(y)
    if y.dtype == np.int16:
        return np.ascontiguousarray(y)
    return np.ascontiguousarray(y, dtype=np.float32)


def spectrograms_from_waves(waves, precision=None, do_trim=True, want_mag=True, return_bounds=False):
    """utils.py:248-276 for a list of already-decoded waveforms (float32 in [-1,1) or int16 PCM).

    Returns a list of (mel, mag) with the reference's layout: mel is [T, 80] F-contiguous, mag is
    [T, 1025] C-contiguous (so np.save writes the same .npy bytes as prepare_data.py:33-34)."""
    eng = get_engine()
    precision = precision or DEFAULT_PRECISION
    waves = [_as_wave_array(w) for w in waves]
    if not waves:
        return []
    if len({w.dtype for w in waves}) != 1:
        waves = [w.astype(np.float32) / np.float32(32768.0) if w.dtype == np.int16 else w for w in waves]
    lens = np.array([w.size for w in waves], dtype=np.int64)
    if np.any(lens == 0):
        raise IndexError("index 0 is out of bounds for axis 0 with size 0")      # what utils.py:251 raises on empty audio
    offsets_np = np.concatenate([[0], np.cumsum(lens)])
    wave = torch.from_numpy(np.concatenate(waves)).to(eng.device, non_blocking=False)
    offsets = torch.from_numpy(offsets_np).to(eng.device)
    if do_trim:
        bounds = eng.trim_bounds(wave, offsets, int(lens.max()))
    else:
        bounds = torch.from_numpy(np.stack([np.zeros_like(lens), lens], axis=1).astype(np.int32)).to(eng.device)
    frame_off = eng.frame_offsets(bounds)
    eng.check()
    fo = frame_off.cpu().numpy()
    mel_d, mag_d = eng.frontend(wave, offsets, bounds, frame_off, int(fo[-1]), precision=precision, want_mag=want_mag)
    mel_h = mel_d.cpu().numpy()
    mag_h = mag_d.cpu().numpy() if want_mag else None
    bounds_h = bounds.cpu().numpy() if return_bounds else None
    out = []
    for b in range(len(waves)):
        T = int(fo[b + 1] - fo[b])
        mel = mel_h[NMELS * fo[b]:NMELS * fo[b + 1]].reshape(NMELS, T).T        # [T, 80], F-contiguous like the reference
        mag = mag_h[fo[b]:fo[b + 1]] if want_mag else None
        out.append((mel, mag, tuple(int(x) for x in bounds_h[b])) if return_bounds else (mel, mag))
    return out


def get_spectrograms_batch(fpaths, precision=None):
    return spectrograms_from_waves([read_wav(p)[0] for p in fpaths], precision=precision)


def get_spectrograms(fpath):
    '''Parse the wave file in `fpath` and
    Returns normalized melspectrogram and linear spectrogram.
    Args:
      fpath: A string. The full path of a sound file.
    Returns:
      mel: A 2d array of shape (T, n_mels) and dtype of float32.
      mag: A 2d array of shape (T, 1+n_fft/2) and dtype of float32.
    '''
    y, _ = read_wav(fpath)
    return spectrograms_from_waves([y])[0]


def _gl_batch(mags, n_iter, momentum, angles0, power, deemphasis, do_trim):
    eng = get_engine()
    mags = [np.ascontiguousarray(m, dtype=np.float32) for m in mags]
    for m in mags:
        if m.ndim != 2 or m.shape[1] != NBINS:
            raise ValueError("expected [T, %d] magnitude spectrograms, got %s" % (NBINS, (m.shape,)))
    frames = [m.shape[0] for m in mags]
    mag_d = torch.from_numpy(np.concatenate(mags, axis=0)).to(eng.device)
    a0_d = None
    if angles0 is not None:
        a0 = np.concatenate([np.ascontiguousarray(a, dtype=np.complex64) for a in angles0], axis=0)
        a0_d = torch.from_numpy(a0).to(eng.device)
    wav_d, wav_off, bounds_d = eng.griffinlim(mag_d, frames, n_iter=n_iter, momentum=momentum, angles0=a0_d,
                                              power=power, deemphasis=deemphasis, want_trim=do_trim)
    eng.check()
    wav_h = wav_d.cpu().numpy()
    bounds_h = bounds_d.cpu().numpy() if do_trim else None
    out = []
    for b in range(len(mags)):
        w = wav_h[wav_off[b]:wav_off[b + 1]]
        if do_trim:
            w = w[bounds_h[b, 0]:bounds_h[b, 1]]
        out.append(np.ascontiguousarray(w, dtype=np.float32))
    return out


def spectrogram2wav_batch(mags, n_iter=None, momentum=0.0, angles0=None, do_trim=True):
    """utils.py:281-306 for a list of [T_b, 1025] normalised magnitude spectrograms.
    angles0: optional list of [T_b, 1025] complex unit phases (frame-major) replacing the zero-phase start."""
    return _gl_batch(mags, ap.n_iter if n_iter is None else n_iter, momentum, angles0, ap.power, True, do_trim)


def spectrogram2wav(mag):
    '''# Generate wave file from linear magnitude spectrogram using Griffin-Lim
    Args:
      mag: A numpy array of (T, 1+n_fft//2)
    Returns:
      wav: A 1-D numpy array.
    '''
    return spectrogram2wav_batch([mag])[0]


def griffin_lim(spectrogram):
    '''Applies Griffin-Lim's raw.  spectrogram: [1+n_fft//2, t] linear magnitude (utils.py:309-320).'''
    S = np.ascontiguousarray(np.asarray(spectrogram, dtype=np.float32).T)
    return _gl_batch([S], ap.n_iter, 0.0, None, 0.0, False, False)[0]


def invert_spectrogram(spectrogram):
    '''Applies inverse fft (librosa.istft, utils.py:323-328).
    Args:
      spectrogram: [1+n_fft//2, t] complex
    '''
    X = np.asarray(spectrogram).T
    S = np.ascontiguousarray(np.abs(X), dtype=np.float32)
    ang = np.where(S > 0, X / np.maximum(S, np.finfo(np.float32).tiny), 1.0).astype(np.complex64)
    return _gl_batch([S], 0, 0.0, [ang], 0.0, False, False)[0]
