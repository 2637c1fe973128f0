"""Masking ops on the padded mel batch: drop-ins for ``utils.specaugment`` and ``utils.noise_fn``.

Same names, arguments, return values and random streams as the reference (src/utils.py:40-75; callers
src/train.py:236, src/network.py:205,307): the draws are made here, on the host side of the boundary, by the very calls
the reference makes and in its order -- ``torch.randint`` twice, then ``random.randrange`` per utterance;
``torch.bernoulli`` on the input's device -- so a seeded run masks the same rows.  The arithmetic (slab mean, row fill,
row scaling) is one pass of ``unast_mel_augment`` (csrc/augment.cuh) over the batch in HBM.

Both spans of the reference run along the TIME axis: ``res[i][:][f_zero:f_zero+f]`` takes a full view and then slices
rows, so the "frequency" mask is a second time mask of up to ``freq_mask`` frames.  That is reproduced, not corrected.
"""
from __future__ import annotations

import random

import numpy as np
import torch

from .engine import get_engine


def draw_specaugment_spans(mel_len, freq_mask=20, time_mask=100):
    """The reference's draws (src/utils.py:55-68) -> int32 [B, 4] rows (f_zero, f, t_zero, t).  Raises what the
    reference raises (``ValueError: empty range`` from ``random.randrange`` when an utterance is shorter than a span)."""
    lens = [int(v) for v in (mel_len.tolist() if hasattr(mel_len, "tolist") else mel_len)]
    shape = tuple(mel_len.shape) if hasattr(mel_len, "shape") else (len(lens),)
    freq_len = torch.randint(0, freq_mask, shape)
    time_len = torch.randint(0, time_mask, shape)
    spans = np.zeros((len(lens), 4), dtype=np.int32)
    for i, n in enumerate(lens):
        f = freq_len[i].item()
        t = time_len[i].item()
        if n - t <= 0:
            t = random.randrange(0, n // 2)
        f_zero = random.randrange(0, n - f)
        t_zero = random.randrange(0, n - t)
        spans[i] = (f_zero, f, t_zero, t)
    return spans


def specaugment(mel, mel_len, freq_mask=20, time_mask=100, replace_with_zero=False):
    """src/utils.py:51-75 for a float32 [B, T, 80] device batch; returns a new tensor (the input is not modified)."""
    eng = get_engine()
    spans = draw_specaugment_spans(mel_len, freq_mask, time_mask)
    x = mel.detach()
    if x.device != eng.device:
        x = x.to(eng.device)
    return eng.mel_augment(x.contiguous(), spans=torch.from_numpy(spans).to(eng.device), replace_with_zero=replace_with_zero)


def noise_fn(to_noise, mask_p=.3, swap_p=0):
    """src/utils.py:40-49: whole time steps of a [batch, seq_len, dim] tensor zeroed with probability ``mask_p``
    (``swap_p`` does nothing in the reference either)."""
    eng = get_engine()
    gen = torch.zeros((to_noise.shape[0], to_noise.shape[1]), device=to_noise.device)
    gen.fill_(1 - mask_p)
    zero_mask = torch.bernoulli(gen)
    x = to_noise.detach()
    if x.device != eng.device:
        x, zero_mask = x.to(eng.device), zero_mask.to(eng.device)
    return eng.mel_augment(x.contiguous(), keep=zero_mask.contiguous())


def specaugment_noise(mel, mel_len, freq_mask=20, time_mask=100, replace_with_zero=False, mask_p=.3):
    """``noise_fn(specaugment(mel, mel_len))`` in a single pass over the batch (draw order: specaugment's, then
    noise_fn's)."""
    eng = get_engine()
    spans = draw_specaugment_spans(mel_len, freq_mask, time_mask)
    gen = torch.zeros((mel.shape[0], mel.shape[1]), device=mel.device)
    gen.fill_(1 - mask_p)
    zero_mask = torch.bernoulli(gen)
    x = mel.detach()
    if x.device != eng.device:
        x, zero_mask = x.to(eng.device), zero_mask.to(eng.device)
    return eng.mel_augment(x.contiguous(), spans=torch.from_numpy(spans).to(eng.device), keep=zero_mask.contiguous(),
                           replace_with_zero=replace_with_zero)


__all__ = ["specaugment", "noise_fn", "specaugment_noise", "draw_specaugment_spans"]
