"""Seeded synthetic LJSpeech-shaped waveforms.

No dataset can be downloaded here, so every benchmark / parity input is generated:
a speech-like harmonic signal (f0 120 Hz +- 30 Hz vibrato, 39 harmonics shaped by two
formant-like resonances, 3 Hz syllabic envelope, peak 0.6) over a white noise floor of
1e-3, with 0.25 s of near-silence (1e-5) before a hard onset and after a hard offset so
that ``librosa.effects.trim`` (reference src/utils.py:248) has something to remove.

The trim decision is discrete (a frame on the -60 dB line moves the cut by 512 samples),
so the generator re-seeds until every trim frame is at least ``margin_db`` away from the
threshold: parity on T / start / end is then exact by construction, not by luck.
"""
from __future__ import annotations

import numpy as np

SR = 22050


def _trim_db(y):
    """Per-frame dB relative to the loudest frame, as librosa.effects.trim computes it."""
    ypad = np.pad(y.astype(np.float64), 1024, mode="reflect")
    n_frames = 1 + len(y) // 512
    cs = np.concatenate([[0.0], np.cumsum(ypad * ypad)])
    starts = 512 * np.arange(n_frames)
    mse = (cs[starts + 2048] - cs[starts]) / 2048.0
    db = 10.0 * np.log10(np.maximum(1e-10, mse))
    return db - db.max()


def speech_like(duration_s, seed, margin_db=3.0, lead_s=0.25):
    """float32 waveform in [-1, 1) of round(duration_s*SR) samples."""
    n = int(round(duration_s * SR))
    for attempt in range(64):
        rng = np.random.RandomState(seed + 7919 * attempt)
        t = np.arange(n, dtype=np.float64) / SR
        f0 = 120.0 + 30.0 * np.sin(2 * np.pi * (0.7 + 0.3 * rng.rand()) * t + 2 * np.pi * rng.rand())
        phase = 2 * np.pi * np.cumsum(f0) / SR
        f1, f2 = 500.0 + 300.0 * rng.rand(), 1500.0 + 900.0 * rng.rand()
        y = np.zeros(n)
        for h in range(1, 40):
            fh = 120.0 * h
            gain = 1.0 / (1.0 + ((fh - f1) / 180.0) ** 2) + 0.5 / (1.0 + ((fh - f2) / 300.0) ** 2) + 0.02
            y += gain * np.sin(h * phase + 2 * np.pi * rng.rand())
        env = 0.55 + 0.45 * np.sin(2 * np.pi * 3.0 * t + 2 * np.pi * rng.rand())
        y *= env
        y *= 0.6 / np.max(np.abs(y))
        y += 1e-3 * rng.randn(n)
        lead = min(int(lead_s * SR), n // 8)
        if lead > 0:
            y[:lead] = 1e-5 * rng.randn(lead)
            y[n - lead:] = 1e-5 * rng.randn(lead)
        y = np.clip(y, -1.0, 1.0 - 2.0 ** -15).astype(np.float32)
        db = _trim_db(y)
        if np.all(np.abs(db + 60.0) >= margin_db):
            return y
    raise RuntimeError("could not generate a waveform with a %g dB trim margin" % margin_db)


def to_pcm16(y):
    """Quantise like a 16-bit WAV writer would; librosa.load gives back q/32768."""
    return np.clip(np.round(y.astype(np.float64) * 32768.0), -32768, 32767).astype(np.int16)


def ljspeech_durations(count, seed=1):
    """LJSpeech-shaped clip lengths: clip(N(6.57, 2.19), 1, 10) seconds."""
    rng = np.random.RandomState(seed)
    return np.clip(rng.normal(6.57, 2.19, size=count), 1.0, 10.0)


def write_wav_pcm16(path, y, rate=SR):
    import struct
    q = to_pcm16(y) if y.dtype != np.int16 else y
    data = q.astype("<i2").tobytes()
    with open(path, "wb") as fh:
        fh.write(b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVE")
        fh.write(b"fmt " + struct.pack("<IHHIIHH", 16, 1, 1, rate, rate * 2, 2, 16))
        fh.write(b"data" + struct.pack("<I", len(data)) + data)
