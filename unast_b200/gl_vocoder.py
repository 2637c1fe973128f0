"""Drop-in for the reference's ``src/gl_vocoder.py``: ``vocode(list_file, mels_dir, out_dir)``.

Same CLI, same file naming (``<id>.mag.npy`` in, ``<id>.wav`` out, gl_vocoder.py:15,34), same
algorithm as ``librosa.griffinlim(S**power, hop_length, win_length)`` at gl_vocoder.py:26 (32
iterations, momentum 0.99, random initial phase drawn from the global ``np.random`` exactly as
librosa draws it: one ``rand(1025, T)`` per utterance, in list order), but the utterances of the
list are processed as GPU batches instead of the reference's serial Python loop (gl_vocoder.py:14).
"""
from __future__ import annotations

import argparse
import os

import numpy as np

from . import audio_parameters as ap
from .utils import _gl_batch
from .wavio import write_wav_float32, write_wav_pcm16

GL_ITERS = 32        # librosa.griffinlim default n_iter
GL_MOMENTUM = 0.99   # librosa.griffinlim default momentum


def _read_list(list_file):
    names = []
    with open(list_file, "r", encoding="utf-8") as fh:
        for line in fh:
            line = line.rstrip("\n")
            if line:
                names.append(line.split("|")[0])
    return names


def vocode_arrays(mags, n_iter=GL_ITERS, momentum=GL_MOMENTUM, angles0=None, random_state=None, do_trim=True):
    """Body of the loop at gl_vocoder.py:15-32 for a list of [T, 1025] normalised magnitudes."""
    if angles0 is None:
        rng = np.random if random_state is None else np.random.RandomState(seed=random_state)
        angles0 = [np.exp(2j * np.pi * rng.rand(1 + ap.n_fft // 2, m.shape[0])).T for m in mags]
    return _gl_batch(mags, n_iter, momentum, angles0, ap.power, True, do_trim)


def vocode_tensors(mag, frames, n_iter=GL_ITERS, momentum=GL_MOMENTUM, angles0=None, random_state=None, do_trim=True):
    """Device-tensor entry: the body of the loop at gl_vocoder.py:15-32 for predicted magnitudes that
    are already on the GPU -- e.g. the output of the CBHG `Vocoder` (src/network.py:627-655) as `inf_vocoder.make_mags`
    produces it (src/inf_vocoder.py:48-64) -- so `inf_vocoder.py` + `gl_vocoder.py` become one GPU pipeline with no .npy round
    trip.  mag: float32 device tensor [sum T, 1025] (utterances back to back) or [B, T_max, 1025] with `frames` = the valid
    length of each row.  Returns (wav float32 device tensor, wav_off int64 numpy [B+1], bounds int32 device [B, 2] or None).
    Initial phases are drawn like librosa draws them (one rand(1025, T) per utterance from the global / seeded numpy RNG)."""
    import torch
    from .engine import get_engine
    eng = get_engine()
    frames = [int(t) for t in frames]
    if mag.dim() == 3:                                                       # padded batch -> utterances back to back
        mag = torch.cat([mag[b, :t] for b, t in enumerate(frames)], dim=0)
    mag = mag.detach().to(device=eng.device, dtype=torch.float32).contiguous()
    a0 = None
    if momentum != 0.0 or angles0 is not None:
        if angles0 is None:
            rng = np.random if random_state is None else np.random.RandomState(seed=random_state)
            angles0 = np.concatenate([np.exp(2j * np.pi * rng.rand(1 + ap.n_fft // 2, t)).T for t in frames], axis=0)
        a0 = torch.as_tensor(np.ascontiguousarray(angles0, dtype=np.complex64)).to(eng.device)
    return eng.griffinlim(mag, frames, n_iter=n_iter, momentum=momentum, angles0=a0, power=ap.power, deemphasis=True, want_trim=do_trim)


def vocode(list_file, mels_dir, out_dir, batch_size=64, subtype="PCM_16"):
    """gl_vocoder.py:10-34.  ``subtype``: "PCM_16" (soundfile's default for .wav, what the reference writes) or "FLOAT".
    A reader thread loads the .mag.npy of the next batch and a writer thread stores the wavs of the previous one while the
    GPU runs the current batch (the reference's loop is serial, gl_vocoder.py:14); the random phases are drawn on the main
    thread, in list order, so the global numpy RNG is consumed exactly as the reference consumes it."""
    from concurrent.futures import ThreadPoolExecutor
    if subtype not in ("PCM_16", "FLOAT"):
        raise ValueError("subtype must be 'PCM_16' or 'FLOAT'")
    write = write_wav_pcm16 if subtype == "PCM_16" else write_wav_float32
    if not os.path.exists(out_dir):
        os.makedirs(out_dir)
    names = _read_list(list_file)
    batches = [names[i:i + batch_size] for i in range(0, len(names), batch_size)]

    def load(chunk):
        return [np.load(os.path.join(mels_dir, "{}.mag.npy".format(n))) for n in chunk]

    def store(chunk, wavs):
        for n, wav in zip(chunk, wavs):
            write(os.path.join(out_dir, "{}.wav".format(n)), wav, ap.sr)

    with ThreadPoolExecutor(max_workers=1) as reader, ThreadPoolExecutor(max_workers=1) as writer:
        nxt = reader.submit(load, batches[0]) if batches else None
        pending = None
        for bi, chunk in enumerate(batches):
            mags = nxt.result()
            nxt = reader.submit(load, batches[bi + 1]) if bi + 1 < len(batches) else None
            wavs = vocode_arrays(mags)
            if pending is not None:
                pending.result()
            pending = writer.submit(store, chunk, wavs)
        if pending is not None:
            pending.result()


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument('--list_file', help='csv file with a list of all the LJ examples to vocode')
    parser.add_argument('--mels_dir', help='dir with all the mels to predict')
    parser.add_argument('--out_dir', help='dir to output the wavs to')
    args = parser.parse_args()
    vocode(args.list_file, args.mels_dir, args.out_dir)
