"""Host-buffer streaming front end: the GPU replacement for the reference's
``DataLoader(PrepareDataset, batch_size=1, num_workers=8)`` loop (src/prepare_data.py:29-46).

The reference computes one utterance per fork worker on the host.  Here the parent process owns the
GPU and streams *chunks* of utterances through three CUDA streams so that PCIe and the kernels overlap:

    copy-in  stream : pinned int16/float32 samples  -> device            (chunk i+1)
    compute  stream : trim bounds -> frame offsets -> fused front end    (chunk i)
    copy-out stream : mel / mag / bounds / frame offsets -> pinned host   (chunk i-1)

Results land in pinned host buffers laid out exactly like the .npy payloads of prepare_data.py:33-34
(mel: [80][T_b] per utterance == fortran_order (T_b, 80); mag: [T_b][1025]), and are handed to a
``consume`` callback (e.g. the .npy writer) before the slot is reused.
"""
from __future__ import annotations

import numpy as np
import torch

from .engine import HOP, NBINS, NMELS, get_engine


def plan_chunks(lengths, max_utts=512, max_samples=64 * 220500):
    """Cut an utterance list into consecutive chunks bounded in utterance count and total samples."""
    chunks, u0, acc = [], 0, 0
    for i, n in enumerate(lengths):
        n = int(n)
        if i > u0 and (i - u0 >= max_utts or acc + n > max_samples):
            chunks.append((u0, i))
            u0, acc = i, 0
        acc += n
    if len(lengths) > u0:
        chunks.append((u0, len(lengths)))
    return chunks


# Pinned host buffers are expensive to create (cudaHostAlloc: ~0.3 s per GB) and identical from call to call for a given
# chunk geometry: FrontendStream.close() parks them here, the next stream of the same geometry takes them back.
_PIN_CACHE = {}


def pinned_empty(shape, dtype):
    shape = tuple(int(v) for v in (shape if isinstance(shape, (tuple, list)) else (shape,)))
    free = _PIN_CACHE.get((shape, dtype))
    if free:
        return free.pop()
    return torch.empty(shape, dtype=dtype).pin_memory()


def pinned_release(t):
    _PIN_CACHE.setdefault((tuple(t.shape), t.dtype), []).append(t)


class _Slot:
    def __init__(self, eng, max_utts, max_samples, max_frames, wave_dtype, want_mag):
        dev = eng.device
        self.wave = torch.empty(max_samples, dtype=wave_dtype, device=dev)
        self.off = torch.empty(max_utts + 1, dtype=torch.int64, device=dev)
        self.mel = torch.empty(NMELS * max_frames, dtype=torch.float32, device=dev)
        self.mag = torch.empty((max_frames, NBINS), dtype=torch.float32, device=dev) if want_mag else None
        self.h_off = pinned_empty(max_utts + 1, torch.int64)
        self.h_mel = pinned_empty(NMELS * max_frames, torch.float32)
        self.h_mag = pinned_empty((max_frames, NBINS), torch.float32) if want_mag else None
        self.h_bounds = pinned_empty((max_utts, 2), torch.int32)
        self.h_fo = pinned_empty(max_utts + 1, torch.int32)
        self.ev_in = torch.cuda.Event()
        self.ev_done = torch.cuda.Event()
        self.ev_meta = torch.cuda.Event()
        self.ev_out = torch.cuda.Event()
        self.bounds = self.fo = None
        self.chunk = None
        self.fut = None                         # writer job that still reads this slot's pinned results


class FrontendStream:
    """Pipelined get_spectrograms over host-resident utterances (pinned memory in, pinned memory out)."""

    def __init__(self, lengths, chunks=None, wave_dtype=torch.int16, precision="f64", want_mag=True, n_slots=3,
                 device=None, geometry=None):
        self.eng = get_engine(device)
        self.lengths = np.asarray(lengths, dtype=np.int64)
        self.offsets = np.concatenate([[0], np.cumsum(self.lengths)])
        self.chunks = chunks if chunks is not None else plan_chunks(self.lengths)
        self.precision = precision
        self.want_mag = want_mag
        mu = max(b - a for a, b in self.chunks)
        ms = max(int(self.offsets[b] - self.offsets[a]) for a, b in self.chunks)
        mf = max(int((1 + self.lengths[a:b] // HOP).sum()) for a, b in self.chunks)     # untrimmed upper bound
        if geometry is not None:                                    # (max utterances, max samples) of ANY chunk plan_chunks can
            mu = max(mu, int(geometry[0]))                          # cut with these bounds: buffers then have the same shapes
            ms = max(ms, int(geometry[1]))                          # from call to call and come out of the pinned cache
            mf = max(mf, ms // HOP + mu)
        self.slots = [_Slot(self.eng, mu, ms, mf, wave_dtype, want_mag) for _ in range(n_slots)]
        # size the engine's scratch once: it is shared by the chunks in flight on the compute stream
        self.eng._workspace("trim", self.eng.lib.unast_trim_workspace_bytes(mu, int(self.lengths.max())))
        self.eng._workspace("fe", self.eng.lib.unast_frontend_workspace_bytes(mu))
        dev = self.eng.device
        self.s_in, self.s_comp, self.s_out = (torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.Stream(dev))
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def close(self):
        """Hand the pinned result buffers to the cache (the stream must not be used afterwards)."""
        for sl in self.slots:
            for t in (sl.h_off, sl.h_mel, sl.h_mag, sl.h_bounds, sl.h_fo):
                if t is not None:
                    pinned_release(t)
        self.slots = []

    # -- pipeline stages -----------------------------------------------------------------------
    def _issue_compute(self, slot, wave_host, chunk, chunk_local=False):
        a, b = chunk
        s0, s1 = int(self.offsets[a]), int(self.offsets[b])
        B = b - a
        if slot.fut is not None:                                    # a writer thread may still be reading the pinned results
            slot.fut.result()
            slot.fut = None
        slot.chunk = chunk
        slot.h_off[:B + 1] = torch.from_numpy(self.offsets[a:b + 1] - s0)
        with torch.cuda.stream(self.s_in):
            self.s_in.wait_event(slot.ev_out)                       # slot's previous results have left the device
            slot.wave[:s1 - s0].copy_(wave_host[:s1 - s0] if chunk_local else wave_host[s0:s1], non_blocking=True)
            slot.off[:B + 1].copy_(slot.h_off[:B + 1], non_blocking=True)
            slot.ev_in.record(self.s_in)
        self.h2d_bytes += (s1 - s0) * wave_host.element_size() + (B + 1) * 8
        with torch.cuda.stream(self.s_comp):
            self.s_comp.wait_event(slot.ev_in)
            off = slot.off[:B + 1]
            slot.bounds = self.eng.trim_bounds(slot.wave, off, int(self.lengths[a:b].max()))
            slot.fo = self.eng.frame_offsets(slot.bounds)
            self.eng.frontend(slot.wave, off, slot.bounds, slot.fo, 0, precision=self.precision, want_mag=self.want_mag,
                              mel_out=slot.mel, mag_out=slot.mag)
            slot.ev_done.record(self.s_comp)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(slot.ev_done)
            slot.h_fo[:B + 1].copy_(slot.fo, non_blocking=True)
            slot.h_bounds[:B].copy_(slot.bounds, non_blocking=True)
            slot.ev_meta.record(self.s_out)
        self.d2h_bytes += (B + 1) * 4 + B * 8

    def _issue_copy_out(self, slot):
        slot.ev_meta.synchronize()                                  # frame counts are data dependent (trim)
        B = slot.chunk[1] - slot.chunk[0]
        F = int(slot.h_fo[B])
        with torch.cuda.stream(self.s_out):
            slot.h_mel[:NMELS * F].copy_(slot.mel[:NMELS * F], non_blocking=True)
            if self.want_mag:
                slot.h_mag[:F].copy_(slot.mag[:F], non_blocking=True)
            slot.ev_out.record(self.s_out)
        self.d2h_bytes += F * (NMELS + (NBINS if self.want_mag else 0)) * 4
        return F

    def run(self, wave_host, consume=None, provider=None, writers=0):
        """wave_host: pinned 1-D tensor holding every utterance back to back -- or None with `provider(i)` returning the
        pinned samples of chunk i alone (called from a reader thread, two chunks ahead: file reads overlap the GPU).
        consume(chunk, frame_off, bounds, mel, mag) is called with host views valid until it returns; with writers > 0 it
        runs on a pool of that many writer threads (file writes overlap the GPU; a slot is reused only after its writer
        has returned).  Returns (total frames, total trimmed samples)."""
        self.eng.check()
        n = len(self.chunks)
        pool = reader = None
        if writers > 0 and consume is not None:
            from concurrent.futures import ThreadPoolExecutor
            pool = ThreadPoolExecutor(max_workers=writers)
        ahead, in_done = {}, {}
        if provider is not None:
            from concurrent.futures import ThreadPoolExecutor
            reader = ThreadPoolExecutor(max_workers=1)
            for j in range(min(2, n)):
                ahead[j] = reader.submit(provider, j)
        frames = samples = 0
        pending = []                                                 # slots whose copy-out has not been issued
        draining = []                                                # slots whose copy-out is in flight

        def finish(slot):
            nonlocal frames, samples
            slot.ev_out.synchronize()
            B = slot.chunk[1] - slot.chunk[0]
            fo = slot.h_fo[:B + 1].numpy()
            bd = slot.h_bounds[:B].numpy()
            frames += int(fo[B])
            samples += int((bd[:, 1] - bd[:, 0]).sum())
            if consume is not None:
                F = int(fo[B])
                a = (slot.chunk, fo.copy(), bd.copy(), slot.h_mel[:NMELS * F].numpy(), slot.h_mag[:F].numpy() if self.want_mag else None)
                if pool is not None:
                    slot.fut = pool.submit(consume, *a)
                else:
                    consume(*a)

        for i in range(n):
            slot = self.slots[i % len(self.slots)]
            while any(s is slot for s in draining):                 # slot still owned by an earlier chunk
                finish(draining.pop(0))
            if provider is not None:
                chunk_wave = ahead.pop(i).result()
                if i + 2 < n:
                    if i - 2 in in_done:                             # a provider may recycle the staging buffer of chunk i - 2
                        in_done.pop(i - 2).synchronize()
                    ahead[i + 2] = reader.submit(provider, i + 2)
                slot.src = chunk_wave                                # keep the pinned source alive until its copy has run
                self._issue_compute(slot, chunk_wave, self.chunks[i], chunk_local=True)
                in_done[i] = torch.cuda.Event()
                in_done[i].record(self.s_in)
            else:
                self._issue_compute(slot, wave_host, self.chunks[i])
            pending.append(slot)
            if len(pending) > 1:                                     # software pipelining by one chunk
                s = pending.pop(0)
                self._issue_copy_out(s)
                draining.append(s)
        while pending:
            s = pending.pop(0)
            self._issue_copy_out(s)
            draining.append(s)
        while draining:
            finish(draining.pop(0))
        for slot in self.slots:
            if slot.fut is not None:
                slot.fut.result()
                slot.fut = None
        if pool is not None:
            pool.shutdown()
        if reader is not None:
            reader.shutdown()
        self.eng.check()
        return frames, samples
