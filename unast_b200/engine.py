"""Batched device-side interface to the spectral kernels.

torch is used for device memory, streams and (elsewhere) torch.distributed only -- every transform
is computed by libunast_spectral.so.  ``SpectralEngine`` is what ``utils.get_spectrograms`` /
``utils.spectrogram2wav`` / ``gl_vocoder.vocode`` / ``prepare_data`` are thin wrappers around; it
works on whole batches because one utterance cannot fill 148 SMs.

Batch layout (DESIGN.md "data layout in HBM"):
  wave      1-D float32 or int16 tensor, utterances back to back; ``offsets`` int64 [B+1]
  bounds    int32 [B, 2]  trim start / end inside each utterance (librosa.effects.trim)
  frame_off int32 [B+1]   prefix sum of T_b = 1 + (end - start) // 275
  mag       float32 [sum T, 1025]      utterance b = rows frame_off[b]:frame_off[b+1]
  mel       float32 [80 * sum T]       utterance b = [80, T_b] block at 80 * frame_off[b]
            (or [B, T_max, 80] zero padded for the loader path)
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib as L
from . import audio_parameters as ap

HOP, WIN, NBINS, NMELS, NFFT = ap.hop_length, ap.win_length, 1 + ap.n_fft // 2, ap.n_mels, ap.n_fft


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


class SpectralEngine:
    def __init__(self, device=None):
        if not torch.cuda.is_available():
            raise RuntimeError("unast_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.lib = L.load()
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        h = C.c_void_p()
        L.check(self.lib.unast_create(self.device.index, C.byref(h)))
        self.h = h
        self._ws = {}
        self.num_sms = self.lib.unast_num_sms(self.h)
        self.launches = 0                      # kernels of ours launched through this engine (bench "gpu_launches")

    def __del__(self):
        try:
            if getattr(self, "h", None):
                self.lib.unast_destroy(self.h)
                self.h = None
        except Exception:
            pass

    # -- helpers -------------------------------------------------------------------------------
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _workspace(self, key, nbytes, zero=False):
        t = self._ws.get(key)
        if t is None or t.numel() < nbytes:
            t = (torch.zeros if zero else torch.empty)(max(int(nbytes), 256), dtype=torch.uint8, device=self.device)
            self._ws[key] = t
        return t

    def check(self):
        """Synchronise and raise if a kernel flagged a data error (too-short / empty utterance)."""
        L.check(self.lib.unast_check(self.h, self._stream()))

    def mel_basis(self):
        out = np.empty((NMELS, NBINS), np.float32)
        L.check(self.lib.unast_get_mel_basis(self.h, out.ctypes.data_as(C.c_void_p)))
        return out

    @staticmethod
    def _dtype_code(wave):
        if wave.dtype == torch.float32:
            return L.F32
        if wave.dtype == torch.int16:
            return L.I16
        raise TypeError("wave must be float32 or int16, got %s" % wave.dtype)

    # -- forward path ----------------------------------------------------------------------------
    def trim_bounds(self, wave, offsets, max_len):
        """librosa.effects.trim bounds for every utterance -> int32 [B, 2] (device)."""
        B = offsets.numel() - 1
        bounds = torch.empty((B, 2), dtype=torch.int32, device=self.device)
        nbytes = self.lib.unast_trim_workspace_bytes(B, int(max_len))
        ws = self._workspace("trim", nbytes)
        L.check(self.lib.unast_trim_bounds(self.h, _ptr(wave), self._dtype_code(wave), _ptr(offsets), B, int(max_len),
                                           _ptr(bounds), _ptr(ws), nbytes, self._stream()))
        self.launches += 2
        return bounds

    def frame_offsets(self, bounds):
        B = bounds.shape[0]
        fo = torch.empty(B + 1, dtype=torch.int32, device=self.device)
        L.check(self.lib.unast_frame_offsets(self.h, _ptr(bounds), B, _ptr(fo), self._stream()))
        self.launches += 1
        return fo

    def frontend(self, wave, offsets, bounds, frame_off, total_frames, precision="f64", want_mag=True,
                 mel_layout="per_utt", T_max=0, mel_out=None, mag_out=None):
        """utils.py:251-276 for a whole batch.  Returns (mel, mag) device tensors (layouts above)."""
        B = offsets.numel() - 1
        prec = {"f32": L.PREC_F32, "f64": L.PREC_F64}[precision]
        if mel_layout == "per_utt":
            layout = L.MEL_MAJOR_PER_UTT
            if mel_out is None:
                mel_out = torch.empty(NMELS * total_frames, dtype=torch.float32, device=self.device)
        elif mel_layout == "padded":
            layout = L.MEL_PADDED_BTM
            if mel_out is None:
                mel_out = torch.zeros((B, T_max, NMELS), dtype=torch.float32, device=self.device)
        else:
            raise ValueError("mel_layout must be 'per_utt' or 'padded'")
        if want_mag and mag_out is None:
            mag_out = torch.empty((total_frames, NBINS), dtype=torch.float32, device=self.device)
        nbytes = self.lib.unast_frontend_workspace_bytes(B)
        ws = self._workspace("fe", nbytes)
        L.check(self.lib.unast_frontend(self.h, _ptr(wave), self._dtype_code(wave), _ptr(offsets), _ptr(bounds), _ptr(frame_off),
                                        B, prec, _ptr(mel_out), layout, int(T_max), _ptr(mag_out if want_mag else None),
                                        _ptr(ws), nbytes, self._stream()))
        self.launches += 2
        return mel_out, (mag_out if want_mag else None)

    # -- inverse path ----------------------------------------------------------------------------
    def griffinlim(self, mag_norm, frames, n_iter=ap.n_iter, momentum=0.0, angles0=None, power=ap.power,
                   deemphasis=True, want_trim=True, wav_out=None):
        """utils.py:289-306 / gl_vocoder.py:17-32 for a batch of normalised magnitude spectrograms.

        mag_norm: float32 [sum T, 1025] device tensor; frames: list of T_b.
        angles0:  complex64 [sum T, 1025] device tensor (unit phases) or None for the zero-phase start.
        Returns (wav float32 [sum 275 (T_b-1)], wav_off numpy int64 [B+1], bounds int32 [B,2] device or None).
        """
        frames_np = np.ascontiguousarray(np.asarray(frames, dtype=np.int32))
        B = int(frames_np.size)
        if mag_norm.dtype != torch.float32 or not mag_norm.is_contiguous():
            raise TypeError("mag_norm must be a contiguous float32 tensor")
        if mag_norm.numel() != int(frames_np.sum()) * NBINS:
            raise ValueError("mag_norm does not match frames")
        a0 = None
        if angles0 is not None:
            if angles0.dtype != torch.complex64 or not angles0.is_contiguous() or angles0.numel() * 1 != int(frames_np.sum()) * NBINS:
                raise TypeError("angles0 must be a contiguous complex64 [sum T, 1025] tensor")
            a0 = torch.view_as_real(angles0)
        fptr = frames_np.ctypes.data_as(C.c_void_p)
        nbytes = self.lib.unast_griffinlim_workspace_bytes(fptr, B, float(momentum))
        if nbytes == 0:
            raise ValueError("every spectrogram needs at least 5 frames (librosa reflect padding)")
        ws = self._workspace("gl", nbytes)
        wav_off = np.concatenate([[0], np.cumsum(HOP * (frames_np.astype(np.int64) - 1))])
        if wav_out is None:
            wav_out = torch.empty(int(wav_off[-1]), dtype=torch.float32, device=self.device)
        bounds = torch.empty((B, 2), dtype=torch.int32, device=self.device) if want_trim else None
        L.check(self.lib.unast_griffinlim(self.h, _ptr(mag_norm), fptr, B, int(n_iter), float(momentum), _ptr(a0), float(power),
                                          1 if deemphasis else 0, _ptr(wav_out), _ptr(bounds), _ptr(ws), nbytes, self._stream()))
        # prepare, (zero-phase start: the float64 first inverse,) ONE persistent Griffin-Lim launch, de-emphasis (+ trim)
        self.launches += 3 + (1 if angles0 is None else 0) + (2 if want_trim else 0)
        return wav_out, wav_off, bounds

    # -- sharded long-form Griffin-Lim ------------------------------------------
    def gl_layout(self, frames, momentum=0.0):
        """Byte offsets (ola0, ola1, ola2, mailboxes, total) of a Griffin-Lim workspace for these fragment sizes."""
        fr = np.ascontiguousarray(np.asarray(frames, dtype=np.int32))
        out = (C.c_size_t * 5)()
        L.check(self.lib.unast_griffinlim_layout(fr.ctypes.data_as(C.c_void_p), int(fr.size), float(momentum), C.cast(out, C.c_void_p)))
        return [int(v) for v in out]

    @staticmethod
    def gl_pad_offsets(frames):
        """pad_off of the C ABI: prefix sum of 2048 + 275 (n - 1) floats over the fragments."""
        fr = np.asarray(frames, dtype=np.int64)
        return np.concatenate([[0], np.cumsum(NFFT + HOP * (fr - 1))])

    def fragment_samples(self, t0, n, tglob):
        a, b = C.c_int64(), C.c_int64()
        L.check(self.lib.unast_griffinlim_fragment_samples(int(t0), int(n), int(tglob), C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def griffinlim_fragments(self, mag_norm, frames, t0, tglob, links=(), epoch=1, n_iter=ap.n_iter, momentum=0.0, angles0=None,
                             power=ap.power, workspace=None, raw_out=None):
        """Griffin-Lim iterations on fragments [t0, t0 + n) of longer utterances; `links` = dicts with keys fragment, side,
        peer_ola (3 addresses), peer_mail (address), my_mail (slot) -- see include/unast_spectral.h.  Returns
        (raw float32 [sum owned samples], raw_off numpy int64 [B+1]): normalised, not de-emphasised."""
        fr = np.ascontiguousarray(np.asarray(frames, dtype=np.int32))
        t0a = np.ascontiguousarray(np.asarray(t0, dtype=np.int32))
        tga = np.ascontiguousarray(np.asarray(tglob, dtype=np.int32))
        B = int(fr.size)
        if mag_norm.dtype != torch.float32 or not mag_norm.is_contiguous() or mag_norm.numel() != int(fr.sum()) * NBINS:
            raise TypeError("mag_norm must be a contiguous float32 [sum n, 1025] tensor")
        a0 = None
        if angles0 is not None:
            if angles0.dtype != torch.complex64 or not angles0.is_contiguous() or angles0.numel() != int(fr.sum()) * NBINS:
                raise TypeError("angles0 must be a contiguous complex64 [sum n, 1025] tensor")
            a0 = torch.view_as_real(angles0)
        nbytes = self.lib.unast_griffinlim_workspace_bytes(fr.ctypes.data_as(C.c_void_p), B, float(momentum))
        if nbytes == 0:
            raise ValueError("bad fragment sizes")
        if workspace is None:
            ws_ptr = _ptr(self._workspace("glf", nbytes, zero=True))
        else:
            ws_ptr = C.c_void_p(int(workspace))
        counts = [self.fragment_samples(a, n, t) for a, n, t in zip(t0a, fr, tga)]
        raw_off = np.concatenate([[0], np.cumsum([e - s for s, e in counts])]).astype(np.int64)
        if raw_out is None:
            raw_out = torch.empty(int(raw_off[-1]), dtype=torch.float32, device=self.device)
        arr = (L.GlLink * max(1, len(links)))()
        for i, lk in enumerate(links):
            arr[i].fragment, arr[i].side, arr[i].my_mail = int(lk["fragment"]), int(lk["side"]), int(lk["my_mail"])
            for k in range(3):
                arr[i].peer_ola[k] = int(lk["peer_ola"][k])
            arr[i].peer_mail = int(lk["peer_mail"])
        L.check(self.lib.unast_griffinlim_fragments(self.h, _ptr(mag_norm), fr.ctypes.data_as(C.c_void_p), t0a.ctypes.data_as(C.c_void_p),
                                                    tga.ctypes.data_as(C.c_void_p), B, C.cast(arr, C.c_void_p), len(links), int(epoch),
                                                    int(n_iter), float(momentum), _ptr(a0), float(power), _ptr(raw_out), ws_ptr, nbytes,
                                                    self._stream()))
        self.launches += 3 + (1 if angles0 is None else 0)
        return raw_out, raw_off

    def deemph_trim(self, x, offsets, max_len, deemphasis=True, want_trim=True):
        """lfilter([1], [1, -0.97]) in float64 + librosa.effects.trim bounds of plain waveforms (device float32 / int64)."""
        B = offsets.numel() - 1
        out = torch.empty_like(x)
        bounds = torch.empty((B, 2), dtype=torch.int32, device=self.device) if want_trim else None
        nbytes = self.lib.unast_deemph_trim_workspace_bytes(B, int(max_len))
        ws = self._workspace("det", nbytes)
        L.check(self.lib.unast_deemph_trim(self.h, _ptr(x), _ptr(offsets), B, int(max_len), 1 if deemphasis else 0, _ptr(out), _ptr(bounds),
                                           _ptr(ws), nbytes, self._stream()))
        self.launches += 1 + (2 if want_trim else 0)
        return out, bounds

    def peer_alloc(self, nbytes):
        """(device address, 64-byte IPC handle) of zero-filled device memory other processes of the node can map."""
        p, h = C.c_void_p(), (C.c_ubyte * 64)()
        L.check(self.lib.unast_peer_alloc(self.h, int(nbytes), C.byref(p), C.cast(h, C.c_void_p)))
        return int(p.value), bytes(h)

    def peer_open(self, handle):
        p, h = C.c_void_p(), (C.c_ubyte * 64).from_buffer_copy(handle)
        L.check(self.lib.unast_peer_open(self.h, C.cast(h, C.c_void_p), C.byref(p)))
        return int(p.value)

    def peer_close(self, addr):
        L.check(self.lib.unast_peer_close(self.h, C.c_void_p(int(addr))))

    def peer_free(self, addr):
        L.check(self.lib.unast_peer_free(self.h, C.c_void_p(int(addr))))

    def mel_augment(self, x, spans=None, keep=None, replace_with_zero=False, out=None):
        """utils.py:40-75 on a float32 [B, T, D] device batch: row spans (int32 [B, 4] = f_zero, f, t_zero, t) filled
        with the slab mean (or 0) and rows scaled by `keep` (float32 [B, T]); one pass, see csrc/augment.cuh."""
        if x.dtype != torch.float32 or not x.is_contiguous() or x.dim() != 3:
            raise TypeError("x must be a contiguous float32 [B, T, D] tensor")
        B, T, D = (int(v) for v in x.shape)
        if spans is not None and (spans.dtype != torch.int32 or tuple(spans.shape) != (B, 4) or not spans.is_contiguous()):
            raise TypeError("spans must be a contiguous int32 [B, 4] tensor")
        if keep is not None and (keep.dtype != torch.float32 or tuple(keep.shape) != (B, T) or not keep.is_contiguous()):
            raise TypeError("keep must be a contiguous float32 [B, T] tensor")
        if out is None:
            out = torch.empty_like(x)
        nbytes = self.lib.unast_mel_augment_workspace_bytes(B)
        ws = self._workspace("aug", nbytes)
        L.check(self.lib.unast_mel_augment(self.h, _ptr(x), B, T, D, _ptr(spans), _ptr(keep), 1 if replace_with_zero else 0,
                                           _ptr(out), _ptr(ws), nbytes, self._stream()))
        self.launches += 2 if (spans is not None and not replace_with_zero) else 1
        return out

    def mel_augment_static(self, x, out, spans=None, keep=None, replace_with_zero=False):
        """mel_augment for callers that must not allocate (include/unast_spectral.h: unast_mel_augment_static): `out` is
        required, the scratch is the library's file-scope device array (B <= 1024), no torch or CUDA allocation happens here.
        Calls that overlap on the GPU share that array and race."""
        if x.dtype != torch.float32 or not x.is_contiguous() or x.dim() != 3 or out.dtype != torch.float32 or not out.is_contiguous() \
                or out.shape != x.shape:
            raise TypeError("x and out must be contiguous float32 [B, T, D] tensors of the same shape")
        B, T, D = (int(v) for v in x.shape)
        if spans is not None and (spans.dtype != torch.int32 or tuple(spans.shape) != (B, 4) or not spans.is_contiguous()):
            raise TypeError("spans must be a contiguous int32 [B, 4] tensor")
        if keep is not None and (keep.dtype != torch.float32 or tuple(keep.shape) != (B, T) or not keep.is_contiguous()):
            raise TypeError("keep must be a contiguous float32 [B, T] tensor")
        L.check(self.lib.unast_mel_augment_static(self.h, _ptr(x), B, T, D, _ptr(spans), _ptr(keep), 1 if replace_with_zero else 0,
                                                  _ptr(out), self._stream()))
        self.launches += 2 if (spans is not None and not replace_with_zero) else 1
        return out

    def set_gl_block(self, frames):
        """Canonical Griffin-Lim summation blocks (include/unast_spectral.h: unast_griffinlim_set_block): 0 = off (default),
        G >= 4 makes every later Griffin-Lim call of this engine independent of the cut into runs / fragments / ranks."""
        L.check(self.lib.unast_griffinlim_set_block(self.h, int(frames)))
        self.gl_block = int(frames)

    def prenet_pack(self, weight, bias=None):
        """nn.Linear(80 -> n_out) parameters -> the shared-memory image unast_mel_prenet_fc1 uses (repack after an update)."""
        if weight.dtype != torch.float32 or weight.dim() != 2 or weight.shape[1] != NMELS:
            raise TypeError("weight must be a float32 [n_out, 80] tensor")
        n_out = int(weight.shape[0])
        nbytes = self.lib.unast_prenet_packed_bytes(n_out)
        if nbytes == 0:
            raise ValueError("n_out must be a multiple of 32 in 32..256")
        w = weight.detach().to(self.device).contiguous()
        b = bias.detach().to(self.device, torch.float32).contiguous() if bias is not None else None
        packed = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        L.check(self.lib.unast_prenet_pack(self.h, _ptr(w), _ptr(b), n_out, _ptr(packed), self._stream()))
        self.launches += 1
        return packed, n_out

    def mel_prenet_fc1(self, x, packed, n_out, spans=None, keep=None, replace_with_zero=False, relu=True, out=None):
        """relu(fc1((row in a span ? fill : x) * keep)) for a float32 [B, T, 80] device batch -> [B, T, n_out] (csrc/prenet.cuh)."""
        if x.dtype != torch.float32 or not x.is_contiguous() or x.dim() != 3 or x.shape[2] != NMELS:
            raise TypeError("x must be a contiguous float32 [B, T, 80] tensor")
        B, T = int(x.shape[0]), int(x.shape[1])
        if spans is not None and (spans.dtype != torch.int32 or tuple(spans.shape) != (B, 4) or not spans.is_contiguous()):
            raise TypeError("spans must be a contiguous int32 [B, 4] tensor")
        if keep is not None and (keep.dtype != torch.float32 or tuple(keep.shape) != (B, T) or not keep.is_contiguous()):
            raise TypeError("keep must be a contiguous float32 [B, T] tensor")
        if out is None:
            out = torch.empty((B, T, n_out), dtype=torch.float32, device=self.device)
        nbytes = self.lib.unast_mel_augment_workspace_bytes(B)
        ws = self._workspace("aug", nbytes)
        L.check(self.lib.unast_mel_prenet_fc1(self.h, _ptr(x), B, T, _ptr(spans), _ptr(keep), 1 if replace_with_zero else 0, _ptr(packed),
                                              int(n_out), 1 if relu else 0, _ptr(out), _ptr(ws), nbytes, self._stream()))
        self.launches += 2 if (spans is not None and not replace_with_zero) else 1
        return out

    # -- diagnostics -------------------------------------------------------------------------------
    def debug_rfft(self, frames, precision="f32"):
        F = frames.shape[0]
        out = torch.empty((F, NBINS), dtype=torch.complex64, device=self.device)
        L.check(self.lib.unast_debug_rfft(self.h, _ptr(frames), F, {"f32": 0, "f64": 1}[precision],
                                          _ptr(torch.view_as_real(out)), self._stream()))
        return out

    def debug_irfft(self, spec):
        F = spec.shape[0]
        out = torch.empty((F, WIN), dtype=torch.float32, device=self.device)
        L.check(self.lib.unast_debug_irfft(self.h, _ptr(torch.view_as_real(spec)), F, _ptr(out), self._stream()))
        return out


    def pipe_peak_tflops(self, kind="f32", iters=4096, reps=5):
        """Measured sustained FMA rate (TFLOP/s) of the float32 / float64 / packed-float32x2 pipe (CUDA events)."""
        k = {"f32": 0, "f64": 1, "f32x2": 2}[kind]
        out = torch.empty(self.num_sms * 8 * 256, dtype=torch.float32, device=self.device)
        flop = C.c_double()
        best = 0.0
        for r in range(reps + 1):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            L.check(self.lib.unast_debug_pipe_peak(self.h, k, int(iters), _ptr(out), C.byref(flop), self._stream()))
            e1.record()
            e1.synchronize()
            if r:
                best = max(best, flop.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
        return best


_ENGINES = {}


def get_engine(device=None):
    """Process-wide engine per device (tables are built once, like a cuFFT plan)."""
    if not torch.cuda.is_available():
        raise RuntimeError("unast_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    idx = torch.cuda.current_device() if device is None else int(device)
    if idx not in _ENGINES:
        _ENGINES[idx] = SpectralEngine(idx)
    return _ENGINES[idx]
