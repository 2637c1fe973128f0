"""Drop-in for the reference's ``src/prepare_data.py``: every wav of a metadata list ->
``<id>.pt.npy`` (mel, fortran_order (T, 80)) and ``<id>.mag.npy`` (mag, C-order (T, 1025)), written
next to the wav exactly as ``PrepareDataset.__getitem__`` does at prepare_data.py:29-38.

The reference drives one utterance per item through ``DataLoader(batch_size=1, num_workers=8)`` fork
workers (prepare_data.py:42).  A CUDA context cannot live in forked workers, so here the parent
process batches ``batch_size`` utterances per GPU call; with ``torch.distributed`` initialised the list
is dealt longest-first across ranks (no collective on the data path, sharding.py).
"""
from __future__ import annotations

import os

import numpy as np

from .sharding import deal_longest_first
from .utils import spectrograms_from_waves
from .wavio import read_wav


def _read_ids(csv_file):
    ids = []
    with open(csv_file, "r", encoding="utf-8") as fh:
        for line in fh:
            line = line.rstrip("\n")
            if line:
                ids.append(line.split("|")[0])
    return ids


class PrepareDataset:
    """Same constructor / ``__len__`` / ``__getitem__`` contract as prepare_data.py:11-38."""

    def __init__(self, csv_file, root_dir):
        self.ids = _read_ids(csv_file)
        self.root_dir = root_dir

    def load_wav(self, filename):
        return read_wav(filename)

    def __len__(self):
        return len(self.ids)

    def wav_name(self, idx):
        return os.path.join(self.root_dir, self.ids[idx]) + '.wav'

    def __getitem__(self, idx):
        return self.process([idx])[0]

    def process(self, indices, precision=None):
        names = [self.wav_name(i) for i in indices]
        waves = [read_wav(n)[0] for n in names]
        out = []
        for name, (mel, mag) in zip(names, spectrograms_from_waves(waves, precision=precision)):
            np.save(name[:-4] + '.pt', mel)
            np.save(name[:-4] + '.mag', mag)
            out.append({'mel': mel, 'mag': mag})
        return out


def prepare(csv_file, root_dir, batch_size=64, rank=0, world_size=1, precision=None, progress=False):
    """Process this rank's share of the list.  Returns the number of utterances written."""
    ds = PrepareDataset(csv_file, root_dir)
    sizes = [os.path.getsize(ds.wav_name(i)) for i in range(len(ds))]
    mine = deal_longest_first(sizes, world_size)[rank]
    it = range(0, len(mine), batch_size)
    if progress:
        from tqdm import tqdm
        it = tqdm(it)
    for s in it:
        ds.process(mine[s:s + batch_size], precision=precision)
    return len(mine)


if __name__ == '__main__':
    import argparse
    parser = argparse.ArgumentParser()
    parser.add_argument('--metadata', required=True, help="LJSpeech metadata.csv ('|' separated, id first)")
    parser.add_argument('--wavs', required=True, help='directory with the wavs')
    parser.add_argument('--batch_size', type=int, default=64)
    args = parser.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        import torch
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    prepare(args.metadata, args.wavs, args.batch_size, rank, world, progress=(rank == 0))
