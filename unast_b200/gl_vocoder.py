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


def vocode(list_file, mels_dir, out_dir, batch_size=64, subtype="PCM_16"):
    """gl_vocoder.py:10-34.  ``subtype``: "PCM_16" (soundfile's default for .wav, what the reference writes) or "FLOAT"."""
    if subtype not in ("PCM_16", "FLOAT"):
        raise ValueError("subtype must be 'PCM_16' or 'FLOAT'")
    write = write_wav_pcm16 if subtype == "PCM_16" else write_wav_float32
    if not os.path.exists(out_dir):
        os.makedirs(out_dir)
    names = _read_list(list_file)
    for i in range(0, len(names), batch_size):
        chunk = names[i:i + batch_size]
        mags = [np.load(os.path.join(mels_dir, "{}.mag.npy".format(n))) for n in chunk]
        wavs = vocode_arrays(mags)
        for n, wav in zip(chunk, wavs):
            write(os.path.join(out_dir, "{}.wav".format(n)), wav, ap.sr)


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument('--list_file', help='csv file with a list of all the LJ examples to vocode')
    parser.add_argument('--mels_dir', help='dir with all the mels to predict')
    parser.add_argument('--out_dir', help='dir to output the wavs to')
    args = parser.parse_args()
    vocode(args.list_file, args.mels_dir, args.out_dir)
