"""Loader integration: the mel side of the reference's training loader, computed on the fly.

The reference's ``LJDatasets.__getitem__`` (src/preprocess.py:34-51) reads a mel that ``prepare_data.py`` wrote to disk
and ``collate_fn_transformer`` (src/preprocess.py:82-121) sorts the batch by text length (descending), pads text with 0
(``_prepare_data`` :153-155) and mel with 0 to the longest item (``_pad_mel`` :176-182) and returns
``(text int64 [B, L_max], mel float32 [B, T_max, 80], text_length int64 [B], mel_length int64 [B])``.

Here the items carry waveforms instead of mels; one launch of the fused front end writes the sorted, zero-padded
``[B, T_max, 80]`` batch tensor straight into device memory (``UNAST_MEL_PADDED_BTM`` layout of ``unast_frontend``) -- no
.npy round trip, no per-item host work.  Same names, argument meaning, return order and dtypes as the reference; the
tensors live on the GPU (``process_batch``, src/train.py:80-94, moves them there anyway).  The text pipeline itself
(``raw_text_to_phoneme_ids``) is out of scope: items carry their phoneme ids.
"""
from __future__ import annotations

from collections.abc import Mapping

import numpy as np
import torch

from . import audio_parameters as ap
from .engine import get_engine
from .utils import DEFAULT_PRECISION, _as_wave_array


def _pad_data(x, length):
    """src/preprocess.py:149-151"""
    return np.pad(x, (0, length - x.shape[0]), mode="constant", constant_values=0)


def _prepare_data(inputs):
    """src/preprocess.py:153-155"""
    max_len = max((len(x) for x in inputs))
    return np.stack([_pad_data(x, max_len) for x in inputs])


def _sort_by_text_length(batch):
    """The permutation ``sorted(zip(field, text_length), key=lambda x: x[1], reverse=True)`` applies to every field
    (src/preprocess.py:98-104): descending text length, ties in their original order (sorted() is stable, also with
    reverse=True)."""
    text_length = [d["text_length"] if "text_length" in d else len(d["text"]) for d in batch]
    return sorted(range(len(batch)), key=lambda i: text_length[i], reverse=True), text_length


def mel_batch_from_waves(waves, precision=None, do_trim=True, wave_device=None):
    """Waveforms (float32 in [-1, 1) or int16 PCM, one dtype) -> (mel [B, T_max, 80] float32 device tensor, zero padded
    like ``_pad_mel``; mel_length int64 [B] device tensor).  Arithmetic of src/utils.py:248-276, batch order kept."""
    eng = get_engine()
    precision = precision or DEFAULT_PRECISION
    waves = [_as_wave_array(w) for w in waves]
    if not waves:
        raise ValueError("empty batch")
    if len({w.dtype for w in waves}) != 1:
        waves = [w.astype(np.float32) / np.float32(32768.0) if w.dtype == np.int16 else w for w in waves]
    lens = np.array([w.size for w in waves], dtype=np.int64)
    if np.any(lens == 0):
        raise IndexError("index 0 is out of bounds for axis 0 with size 0")      # what utils.py:251 raises on empty audio
    offsets_np = np.concatenate([[0], np.cumsum(lens)])
    host = torch.from_numpy(np.concatenate(waves))
    wave = host.pin_memory().to(eng.device, non_blocking=True) if wave_device is None else wave_device
    offsets = torch.from_numpy(offsets_np).to(eng.device)
    if do_trim:
        bounds = eng.trim_bounds(wave, offsets, int(lens.max()))
    else:
        bounds = torch.from_numpy(np.stack([np.zeros_like(lens), lens], axis=1).astype(np.int32)).to(eng.device)
    frame_off = eng.frame_offsets(bounds)
    eng.check()
    fo = frame_off.cpu().numpy()                                                  # the one host read: T_max sizes the tensor
    frames = np.diff(fo)
    mel, _ = eng.frontend(wave, offsets, bounds, frame_off, int(fo[-1]), precision=precision, want_mag=False,
                          mel_layout="padded", T_max=int(frames.max()))
    return mel, (frame_off[1:] - frame_off[:-1]).to(torch.int64)


def collate_fn_transformer(batch, precision=None):
    """Drop-in for src/preprocess.py:82-121 with on-the-fly mels: items are mappings with ``text`` (phoneme ids),
    ``wave`` (float32 / int16 samples at 22,050 Hz) and optionally ``text_length`` and ``fname``."""
    if not isinstance(batch[0], Mapping):
        raise TypeError("batch must contain tensors, numbers, dicts or lists; found {}".format(type(batch[0])))
    order, text_length = _sort_by_text_length(batch)
    text = _prepare_data([np.asarray(batch[i]["text"]) for i in order]).astype(np.int32)
    mel, mel_length = mel_batch_from_waves([batch[i]["wave"] for i in order], precision=precision)
    dev = mel.device
    out = (torch.as_tensor(text, dtype=torch.long).to(dev), mel,
           torch.as_tensor(sorted(text_length, reverse=True), dtype=torch.long).to(dev), mel_length)
    if "fname" in batch[0]:
        return out, [batch[i]["fname"] for i in order]
    return out


def collate_fn_postnet_mel(batch, precision=None):
    """The mel-only branch of ``collate_fn_postnet`` (src/preprocess.py:123-143): batch order kept, returns the padded
    mel tensor (and ``mel_lens``, ``fnames`` when the items carry ``fname``)."""
    if not isinstance(batch[0], Mapping):
        raise TypeError("batch must contain tensors, numbers, dicts or lists; found {}".format(type(batch[0])))
    mel, mel_length = mel_batch_from_waves([d["wave"] for d in batch], precision=precision)
    if "fname" in batch[0]:
        return mel, [int(x) for x in mel_length.cpu()], [d["fname"] for d in batch]
    return mel


__all__ = ["collate_fn_transformer", "collate_fn_postnet_mel", "mel_batch_from_waves", "_prepare_data", "_pad_data", "ap"]
