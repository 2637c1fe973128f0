"""Long-form front end (case C5): one long waveform, frames sharded over ranks with a read-only halo.

Frames of an STFT are independent, so rank r of `world` takes a contiguous frame range [a, b) of the (globally trimmed)
signal and runs the fused front end on the samples those frames can see: frame t touches signal samples
[275 t - 552, 275 t + 551) (1102 window taps centred on 275 t, plus the previous sample for the pre-emphasis), so a chunk
that starts `HALO` frames early and ends `HALO` frames late contains everything, and the chunk-local frames within
`HALO` of a cut -- the only ones that see the reflect padding / first-sample rule of a chunk boundary that is not a
real signal boundary -- are dropped.  Every kept frame goes through exactly the arithmetic of the single-GPU run
(bit-identical, tests/test_gpu_parity.py); there is no exchange and no collective.

The inverse (Griffin-Lim on one long waveform) would need a seam exchange per iteration and is run as replicas only
(multi-GPU).  Reference semantics: src/utils.py:245-276 (`get_spectrograms`) on a long file.
"""
from __future__ import annotations

import numpy as np
import torch

from . import audio_parameters as ap
from .engine import get_engine
from .utils import DEFAULT_PRECISION, _as_wave_array

HOP = ap.hop_length
HALO = 4                     # frames: 4 * 275 = 1100 >= 552 samples on either side
NMELS = ap.n_mels


def frame_ranges(n_frames, world):
    """Contiguous, near-equal frame ranges [(a, b)] for ranks 0..world-1 (some may be empty for tiny inputs)."""
    edges = [(n_frames * r) // world for r in range(world + 1)]
    return [(edges[r], edges[r + 1]) for r in range(world)]


def chunk_for_range(n_samples, a, b):
    """Sample range [c0, c1) of the trimmed signal that frames [a, b) need (with halo), and the slice of chunk-local
    frames to keep.  c0 is a multiple of the hop, so local frame t' is global frame c0 / 275 + t'."""
    if b <= a:
        return 0, 0, slice(0, 0)
    c0 = max(0, HOP * (a - HALO))
    c1 = min(n_samples, HOP * (b - 1 + HALO) + 1)
    first = a - c0 // HOP
    return c0, c1, slice(first, first + (b - a))


def spectrograms_long(wave, rank=0, world=1, precision=None, do_trim=True):
    """This rank's frames of `get_spectrograms` on one long waveform.

    Returns (mel [Tr, 80] F-contiguous, mag [Tr, 1025], (a, b), T) with Tr = b - a frames of the T global ones."""
    eng = get_engine()
    precision = precision or DEFAULT_PRECISION
    y = _as_wave_array(wave)
    if y.size == 0:
        raise IndexError("index 0 is out of bounds for axis 0 with size 0")
    full = torch.from_numpy(y).to(eng.device)
    if do_trim:                                                            # librosa.effects.trim is a global decision
        off = torch.tensor([0, y.size], dtype=torch.int64, device=eng.device)
        start, end = (int(v) for v in eng.trim_bounds(full, off, int(y.size)).cpu().numpy()[0])
        eng.check()
    else:
        start, end = 0, int(y.size)
    n = end - start
    if n <= 0:
        raise IndexError("index 0 is out of bounds for axis 0 with size 0")
    T = 1 + n // HOP
    a, b = frame_ranges(T, world)[rank]
    c0, c1, keep = chunk_for_range(n, a, b)
    if b <= a:
        return np.zeros((0, NMELS), np.float32, order="F"), np.zeros((0, ap.n_fft // 2 + 1), np.float32), (a, b), T
    chunk = full[start + c0:start + c1].clone()                          # fresh allocation: the kernel wants an aligned base
    nc = c1 - c0
    off = torch.tensor([0, nc], dtype=torch.int64, device=eng.device)
    bounds = torch.tensor([[0, nc]], dtype=torch.int32, device=eng.device)
    frame_off = eng.frame_offsets(bounds)
    eng.check()
    Tc = int(frame_off.cpu().numpy()[-1])
    mel_d, mag_d = eng.frontend(chunk, off, bounds, frame_off, Tc, precision=precision, want_mag=True)
    mel = mel_d.view(NMELS, Tc)[:, keep].contiguous().cpu().numpy().T      # [Tr, 80], F-contiguous like the reference
    mag = mag_d[keep].cpu().numpy()
    return mel, mag, (a, b), T


__all__ = ["frame_ranges", "chunk_for_range", "spectrograms_long", "HALO"]
