"""Long-form front end (case C5): one long waveform, frames sharded over ranks with a read-only halo.

Frames of an STFT are independent, so rank r of `world` takes a contiguous frame range [a, b) of the (globally trimmed)
signal and runs the fused front end on the samples those frames can see: frame t touches signal samples
[275 t - 552, 275 t + 551) (1102 window taps centred on 275 t, plus the previous sample for the pre-emphasis), so a chunk
that starts `HALO` frames early and ends `HALO` frames late contains everything, and the chunk-local frames within
`HALO` of a cut -- the only ones that see the reflect padding / first-sample rule of a chunk boundary that is not a
real signal boundary -- are dropped.  Every kept frame goes through exactly the arithmetic of the single-GPU run
(bit-identical, tests/test_gpu_parity.py); there is no exchange and no collective.

The inverse (Griffin-Lim on one long waveform, src/utils.py:309-320) is sharded the same way -- rank r owns a contiguous
fragment of the frames -- but its iterations are coupled: the 827 samples that the last frames of one fragment share with
the first frames of the next (a *linked seam*) exist in both ranks' overlap-add buffers, and every iteration both sides
add their partial sums into both copies with stores / atomics over NVLink peer mappings (CUDA IPC) and publish their
progress in the neighbour's mailbox.  One persistent kernel per rank and call, no collective and no host round trip
inside the loop (`ShardedGriffinLim`); `griffinlim_fragments_emulated` runs the same protocol with all fragments in
one launch on one GPU (links point into the same workspace), which is how the single-GPU tests cover it.
Reference semantics: src/utils.py:245-276 (`get_spectrograms`), :281-328 (`spectrogram2wav`) on a long file.
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


def frame_ranges(n_frames, world, align=1):
    """Contiguous, near-equal frame ranges [(a, b)] for ranks 0..world-1 (some may be empty for tiny inputs); inner
    edges are multiples of `align`."""
    edges = [(n_frames * r) // world for r in range(world + 1)]
    if align > 1:
        edges = [0] + [min(n_frames, (e + align // 2) // align * align) for e in edges[1:-1]] + [n_frames]
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


SEAM0 = (ap.n_fft - ap.win_length) // 2           # 473: padded position of a frame's first tap relative to 275 t
GL_MIN_FRAGMENT = 9
GL_BLOCK = 4                 # canonical summation block: a sample is covered by at most 5 consecutive frames = 2 blocks


def _links_for(eng, my_frames, my_base, left=None, right=None, momentum=0.0):
    """Link descriptors of ONE fragment (index 0 of this rank's call).  left / right = (partner workspace base address as
    mapped here, partner's fragment size) or None.  Mailbox slots: 0 = my left seam, 1 = my right seam."""
    links = []
    if left is not None:
        base, n_l = left
        lay = eng.gl_layout([n_l], momentum)
        links.append({"fragment": 0, "side": 0, "my_mail": 0, "peer_mail": base + lay[3] + 4 * 1,
                      "peer_ola": [base + lay[k] + 4 * (HOP * n_l + SEAM0) for k in range(3)]})
    if right is not None:
        base, n_r = right
        lay = eng.gl_layout([n_r], momentum)
        links.append({"fragment": 0, "side": 1, "my_mail": 1, "peer_mail": base + lay[3] + 4 * 0,
                      "peer_ola": [base + lay[k] + 4 * SEAM0 for k in range(3)]})
    return links


class ShardedGriffinLim:
    """Griffin-Lim on ONE long spectrogram of T frames, frames sharded by contiguous ranges over the ranks of the default
    process group (one process per GPU).  Construct once per (T, momentum) on every rank (collective: exchanges the CUDA
    IPC handles of the workspaces), then `run()` any number of times.

    canonical=True (default): the overlap-add sums are grouped in blocks of GL_BLOCK frames of the waveform whatever the
    sharding, so the result is bit-identical for every world size (one GPU included), at 10-18 % of the speed.  False: runs
    of equal cost with no block grid -- faster, but the rounding then depends on the partition, and the zero-phase start of
    utils.griffin_lim amplifies that to 60-83 dB of the waveform after 100 iterations."""

    def __init__(self, T, rank, world, momentum=0.0, canonical=True):
        import torch.distributed as dist
        self.eng = get_engine()
        self.T, self.rank, self.world, self.momentum = int(T), int(rank), int(world), float(momentum)
        self.block = GL_BLOCK if canonical else 0
        self.ranges = frame_ranges(self.T, self.world, align=max(1, self.block))
        self.a, self.b = self.ranges[self.rank]
        sizes = [e - s for s, e in self.ranges]
        if self.world > 1 and min(sizes) < GL_MIN_FRAGMENT:
            raise ValueError("every rank needs at least %d frames" % GL_MIN_FRAGMENT)
        lay = self.eng.gl_layout([self.b - self.a], self.momentum)
        self.ws, handle = self.eng.peer_alloc(lay[4])
        self._opened = []
        self.links = []
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, handle)
            left = right = None
            if self.rank > 0:
                base = self.eng.peer_open(handles[self.rank - 1]); self._opened.append(base)
                left = (base, sizes[self.rank - 1])
            if self.rank + 1 < self.world:
                base = self.eng.peer_open(handles[self.rank + 1]); self._opened.append(base)
                right = (base, sizes[self.rank + 1])
            self.links = _links_for(self.eng, self.b - self.a, self.ws, left, right, self.momentum)
            dist.barrier()
        self.epoch = 1

    def run(self, mag_rows, n_iter, angles0_rows=None, power=ap.power, raw_out=None):
        """mag_rows: this rank's rows [a, b) of the normalised magnitude spectrogram (device float32 [b - a, 1025]).
        Returns this rank's owned samples of istft(S * angles) after n_iter iterations (device float32), asynchronous."""
        self.eng.set_gl_block(self.block)
        try:
            raw, _ = self.eng.griffinlim_fragments(mag_rows, [self.b - self.a], [self.a], [self.T], self.links, epoch=self.epoch,
                                                   n_iter=n_iter, momentum=self.momentum, angles0=angles0_rows, power=power,
                                                   workspace=self.ws, raw_out=raw_out)
        finally:
            self.eng.set_gl_block(0)
        self.epoch += int(n_iter) + 2
        return raw

    def sample_range(self):
        return self.eng.fragment_samples(self.a, self.b - self.a, self.T)

    def gather_deemph_trim(self, raw, deemphasis=True, want_trim=True):
        """Epilogue (utils.py:301-306): the ranks' samples are gathered on rank 0 (off the iteration path), de-emphasised in
        float64 and trimmed there.  Returns (wav float32 device tensor, (start, end)) on rank 0, (None, None) elsewhere."""
        import torch.distributed as dist
        L = HOP * (self.T - 1)
        if self.world > 1:
            raw = self.gather_rows(raw)
            if self.rank != 0:
                return None, None
        off = torch.tensor([0, L], dtype=torch.int64, device=raw.device)
        wav, bounds = self.eng.deemph_trim(raw, off, L, deemphasis=deemphasis, want_trim=want_trim)
        if not want_trim:
            return wav, (0, L)
        s, e = (int(v) for v in bounds.cpu().numpy()[0])
        return wav, (s, e)

    def gather_rows(self, x):
        """Concatenation over the ranks (in rank order) of per-rank tensors whose first dimensions differ, on rank 0 (None
        elsewhere).  torch.distributed.gather wants equal shapes: pad to the longest, gather, cut."""
        import torch.distributed as dist
        n = torch.tensor([x.shape[0]], dtype=torch.int64, device=x.device)
        ns = [torch.zeros_like(n) for _ in range(self.world)]
        dist.all_gather(ns, n)
        ns = [int(v.item()) for v in ns]
        m = max(ns)
        pad = torch.zeros((m,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        pad[:x.shape[0]] = x
        parts = [torch.empty_like(pad) for _ in range(self.world)] if self.rank == 0 else None
        dist.gather(pad, parts, dst=0)
        if self.rank != 0:
            return None
        return torch.cat([p[:k] for p, k in zip(parts, ns)])

    def close(self):
        for base in self._opened:
            self.eng.peer_close(base)
        self._opened = []
        if self.ws:
            self.eng.peer_free(self.ws)
            self.ws = 0


_EMU_EPOCH = [1]


def griffinlim_fragments_emulated(mag, world, n_iter, momentum=0.0, angles0=None, power=ap.power, canonical=False):
    """The sharded run's protocol on ONE GPU: the T frames of `mag` (device float32 [T, 1025]) cut into `world` fragments
    that are processed by one launch, every cut a linked seam whose two copies live in the same workspace.  Returns the
    raw (normalised, not de-emphasised) waveform [275 (T - 1)] as a device tensor.  canonical=True: summation blocks of
    GL_BLOCK frames -- the result is then bit-identical for every `world`."""
    eng = get_engine()
    T = int(mag.shape[0])
    ranges = frame_ranges(T, world, align=GL_BLOCK if canonical else 1)
    frames = [e - s for s, e in ranges]
    t0 = [s for s, _ in ranges]
    if world > 8 or (world > 1 and min(frames) < GL_MIN_FRAGMENT):
        raise ValueError("at most 8 fragments of at least %d frames" % GL_MIN_FRAGMENT)
    lay = eng.gl_layout(frames, momentum)
    base = eng._workspace("glf", lay[4], zero=True).data_ptr()
    pad = eng.gl_pad_offsets(frames)
    links = []
    for i in range(world - 1):
        slot_l, slot_r = 2 * i, 2 * i + 1                                  # mailboxes of the left / right side of cut i
        links.append({"fragment": i, "side": 1, "my_mail": slot_l, "peer_mail": base + lay[3] + 4 * slot_r,
                      "peer_ola": [base + lay[k] + 4 * (int(pad[i + 1]) + SEAM0) for k in range(3)]})
        links.append({"fragment": i + 1, "side": 0, "my_mail": slot_r, "peer_mail": base + lay[3] + 4 * slot_l,
                      "peer_ola": [base + lay[k] + 4 * (int(pad[i]) + HOP * frames[i] + SEAM0) for k in range(3)]})
    eng.set_gl_block(GL_BLOCK if canonical else 0)
    try:
        raw, _ = eng.griffinlim_fragments(mag, frames, t0, [T] * world, links, epoch=_EMU_EPOCH[0], n_iter=n_iter, momentum=momentum,
                                          angles0=angles0, power=power, workspace=base)
    finally:
        eng.set_gl_block(0)
    _EMU_EPOCH[0] += int(n_iter) + 2
    return raw


__all__ = ["frame_ranges", "chunk_for_range", "spectrograms_long", "HALO", "ShardedGriffinLim", "griffinlim_fragments_emulated"]
