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


def prepare(csv_file, root_dir, batch_size=128, rank=0, world_size=1, precision=None, progress=False, writers=8):
    """Process this rank's share of the list; returns the number of utterances written.

    The I/O stage as a pipeline: a reader thread parses the wavs of chunk i + 2 into pinned memory
    while chunk i + 1 crosses PCIe, chunk i is on the GPU and chunk i - 1 comes back; `writers` threads emit the .npy files
    straight from the pinned result slots -- numpy's own header bytes, then the payload exactly as it lies in the slot
    (mel [80][T] == fortran_order (T, 80), mag [T][1025]), i.e. the files np.save would write (prepare_data.py:33-34)."""
    import torch
    from .pipeline import FrontendStream, pinned_empty, pinned_release, plan_chunks
    from .wavio import npy_header, wav_info
    ds = PrepareDataset(csv_file, root_dir)
    sizes = [os.path.getsize(ds.wav_name(i)) for i in range(len(ds))]
    mine = deal_longest_first(sizes, world_size)[rank]
    mine = sorted(mine, key=lambda i: (-sizes[i], i))                      # length-bucketed: similar lengths share a chunk
    if not mine:
        return 0
    infos = [wav_info(ds.wav_name(i)) for i in mine]
    lengths = np.array([n for n, _, _ in infos], dtype=np.int64)
    if np.any(lengths == 0):
        raise IndexError("index 0 is out of bounds for axis 0 with size 0")      # what utils.py:251 raises on empty audio
    pcm = all(dt == np.int16 for _, dt, _ in infos)
    wave_dtype = torch.int16 if pcm else torch.float32
    chunks = plan_chunks(lengths, max_utts=batch_size, max_samples=batch_size * 150000)
    offsets = np.concatenate([[0], np.cumsum(lengths)])
    cap = max(batch_size * 150000, int(lengths.max()))                     # any chunk plan_chunks cuts fits: same shapes every call
    stream = FrontendStream(lengths, chunks=chunks, wave_dtype=wave_dtype, precision=precision or "f64", geometry=(batch_size, cap))
    staging = [pinned_empty(cap, wave_dtype) for _ in range(4)]             # chunks i .. i + 3 are alive at most

    bar = None
    if progress:
        from tqdm import tqdm
        bar = tqdm(total=len(mine))

    from concurrent.futures import ThreadPoolExecutor
    wpool = ThreadPoolExecutor(max_workers=max(1, writers))

    def write_range(a, j0, j1, fo, mel, mag):
        for j in range(j0, j1):
            name = ds.wav_name(mine[j])[:-4]
            f0, f1 = int(fo[j - a]), int(fo[j - a + 1])
            T = f1 - f0
            with open(name + '.pt.npy', 'wb') as fh:
                fh.write(npy_header((T, 80), True))
                fh.write(memoryview(mel[80 * f0:80 * f1]))
            with open(name + '.mag.npy', 'wb') as fh:
                fh.write(npy_header((T, 1025), False))
                fh.write(memoryview(mag[f0:f1]))

    def consume(chunk, fo, bounds, mel, mag):
        # the chunk's files are independent: fan them out over the writer threads (write() releases the GIL)
        a, b = chunk
        step = max(1, -(-(b - a) // max(1, writers)))
        jobs = [wpool.submit(write_range, a, j0, min(b, j0 + step), fo, mel, mag) for j0 in range(a, b, step)]
        for job in jobs:
            job.result()
        if bar is not None:
            bar.update(b - a)

    def read_range(view, jobs):
        for j, pos in jobs:
            y = read_wav(ds.wav_name(mine[j]))[0]
            if not pcm and y.dtype == np.int16:
                y = y.astype(np.float32) / np.float32(32768.0)
            view[pos:pos + y.size] = y

    def provider(ci):
        a, b = chunks[ci]
        buf = staging[ci % len(staging)]
        view = buf.numpy()
        pos = offsets[a:b] - offsets[a]
        jobs = [(j, int(pos[j - a])) for j in range(a, b)]
        step = max(1, -(-len(jobs) // max(1, writers)))
        futs = [wpool.submit(read_range, view, jobs[k:k + step]) for k in range(0, len(jobs), step)]
        for f in futs:
            f.result()
        return buf

    try:
        stream.run(None, consume=consume, provider=provider, writers=1)
    finally:
        wpool.shutdown()
        stream.close()
        for t in staging:
            pinned_release(t)
    if bar is not None:
        bar.close()
    return len(mine)


if __name__ == '__main__':
    import argparse
    parser = argparse.ArgumentParser()
    parser.add_argument('--metadata', required=True, help="LJSpeech metadata.csv ('|' separated, id first)")
    parser.add_argument('--wavs', required=True, help='directory with the wavs')
    parser.add_argument('--batch_size', type=int, default=512)
    args = parser.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world > 1:
        import torch
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    prepare(args.metadata, args.wavs, args.batch_size, rank, world, progress=(rank == 0))
