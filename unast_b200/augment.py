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


class _RowMask(torch.autograd.Function):
    """y = x * keep[..., None] through unast_mel_augment, forward and backward (d/dx = keep): the reference's
    ``to_noise * zero_mask`` (src/utils.py:49) is differentiable and is applied to tensors that require grad
    (``embedded_phonemes`` at src/network.py:431,521), so the drop-in must let gradients through."""

    @staticmethod
    def forward(ctx, x, keep):
        ctx.save_for_backward(keep)
        return get_engine().mel_augment(x.detach().contiguous(), keep=keep)

    @staticmethod
    def backward(ctx, grad):
        keep, = ctx.saved_tensors
        return get_engine().mel_augment(grad.contiguous(), keep=keep), None


def noise_fn(to_noise, mask_p=.3, swap_p=0):
    """src/utils.py:40-49 (callers: src/network.py:205,307 on the speech prenet output, :431,:521 on the embedded
    phonemes): whole time steps of a float32 [batch, seq_len, dim] tensor zeroed with probability ``mask_p``
    (``swap_p`` does nothing in the reference either).  Differentiable like the reference's multiply."""
    eng = get_engine()
    if to_noise.dim() != 3 or to_noise.dtype != torch.float32:
        raise TypeError("noise_fn: expected a float32 [batch, seq_len, dim] tensor, got %s %s" % (to_noise.dtype, tuple(to_noise.shape)))
    gen = torch.zeros((to_noise.shape[0], to_noise.shape[1]), device=to_noise.device)
    gen.fill_(1 - mask_p)
    zero_mask = torch.bernoulli(gen)
    x = to_noise
    if x.device != eng.device:
        x, zero_mask = x.to(eng.device), zero_mask.to(eng.device)
    zero_mask = zero_mask.contiguous()
    if x.requires_grad:
        return _RowMask.apply(x, zero_mask)
    return eng.mel_augment(x.contiguous(), keep=zero_mask)


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


class PrenetFC1:
    """specaugment -> noise_fn -> SpeechPrenet.fc1 -> ReLU (src/train.py:236, src/network.py:205, src/module.py:93-95) as ONE
    pass over the padded mel batch, the Linear on the tensor cores (csrc/prenet.cuh).  Inference / frozen-prenet use: the
    kernel has no backward; Dropout (module.py:96) is applied by the caller.  Draw order: specaugment's, then noise_fn's."""

    def __init__(self, weight, bias=None):
        self.eng = get_engine()
        self.packed, self.n_out = self.eng.prenet_pack(weight, bias)

    def repack(self, weight, bias=None):
        self.packed, self.n_out = self.eng.prenet_pack(weight, bias)

    def __call__(self, mel, mel_len=None, augment=True, noise=True, freq_mask=20, time_mask=100, replace_with_zero=False, mask_p=.3,
                 relu=True):
        eng = self.eng
        spans = keep = None
        if augment:
            spans = torch.from_numpy(draw_specaugment_spans(mel_len, freq_mask, time_mask)).to(eng.device)
        if noise:
            gen = torch.zeros((mel.shape[0], mel.shape[1]), device=mel.device)
            gen.fill_(1 - mask_p)
            keep = torch.bernoulli(gen).to(eng.device).contiguous()
        x = mel.detach().to(eng.device).contiguous()
        return eng.mel_prenet_fc1(x, self.packed, self.n_out, spans=spans, keep=keep, replace_with_zero=replace_with_zero, relu=relu)


__all__ = ["specaugment", "noise_fn", "specaugment_noise", "draw_specaugment_spans", "PrenetFC1"]
