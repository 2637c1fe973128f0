"""CPU: host-side logic of the product package -- sharding, WAV/NPY I/O, synthetic data, the C-ABI
library's exported symbols, loud failure without a GPU, and the world_size-2 path over gloo."""
import ctypes
import os
import re
import socket
import subprocess
import sys

import numpy as np
import pytest

from conftest import ROOT
from unast_b200 import audio_parameters as ap
from unast_b200 import sharding, synth, wavio


def test_audio_parameters_match_reference_constants():
    # src/audio_parameters.py:1-20
    assert (ap.sr, ap.n_fft, ap.hop_length, ap.win_length, ap.n_mels) == (22050, 2048, 275, 1102, 80)
    assert (ap.preemphasis, ap.power, ap.max_db, ap.ref_db, ap.n_iter) == (0.97, 1.2, 100, 20, 60)


def test_deal_longest_first_partitions_and_balances():
    sizes = np.round(synth.ljspeech_durations(1000, seed=1) * 80.18).astype(int)
    for world in (1, 2, 4, 8):
        shards = sharding.deal_longest_first(sizes, world)
        flat = sorted(i for s in shards for i in s)
        assert flat == list(range(1000))
        loads = sharding.shard_loads(sizes, shards)
        assert max(loads) - min(loads) <= sizes.max()
        assert max(loads) / (sum(loads) / world) < 1.01


def test_length_buckets_are_sorted_and_complete():
    sizes = [5, 9, 1, 7, 3, 8, 2]
    b = sharding.length_buckets(range(7), sizes, 3)
    assert [len(x) for x in b] == [3, 3, 1]
    flat = [i for x in b for i in x]
    assert [sizes[i] for i in flat] == sorted(sizes, reverse=True)


def test_ragged_and_empty_shards():
    assert sharding.deal_longest_first([], 4) == [[], [], [], []]
    shards = sharding.deal_longest_first([10, 3], 4)
    assert sorted(len(s) for s in shards) == [0, 0, 1, 1]


def test_wav_roundtrip_pcm16_and_float32(tmp_path):
    y = synth.speech_like(1.0, 7)
    p16 = str(tmp_path / "a.wav")
    synth.write_wav_pcm16(p16, y)
    q, rate = wavio.read_wav(p16)
    assert rate == 22050 and q.dtype == np.int16 and np.array_equal(q, synth.to_pcm16(y))
    pf = str(tmp_path / "b.wav")
    wavio.write_wav_float32(pf, y)
    z, rate = wavio.read_wav(pf)
    assert z.dtype == np.float32 and np.array_equal(z, y)
    import scipy.io.wavfile as sw                                                   # third-party reader agrees
    r2, d2 = sw.read(pf)
    assert r2 == 22050 and np.array_equal(d2, y)


def test_vocode_output_format_is_pcm16(tmp_path):
    """gl_vocoder.py:34 `sf.write(path, wav, ap.sr)`: soundfile's default subtype for .wav is PCM_16 (tag 1, 16 bits),
    samples lrint(x * 0x7FFF) with wrap-around (libsndfile, clipping off)."""
    import struct
    y = np.array([0.0, 0.5, -0.5, 1.0, -1.0, 0.25 + 1e-6, 1.00002, 1.5 / 32767.0, 2.5 / 32767.0], np.float64)
    p = str(tmp_path / "v.wav")
    wavio.write_wav_pcm16(p, y)
    b = open(p, "rb").read()
    tag, nch, rate, brate, align, bits = struct.unpack("<HHIIHH", b[20:36])
    assert (tag, nch, rate, brate, align, bits) == (1, 1, 22050, 44100, 2, 16) and b[36:40] == b"data"
    q, _ = wavio.read_wav(p)
    assert q.dtype == np.int16
    assert q.tolist() == [0, 16384, -16384, 32767, -32767, 8192, -32768, 2, 2]    # round half to even; > 1.0 wraps
    import scipy.io.wavfile as sw
    assert np.array_equal(sw.read(p)[1], q)


def test_wav_wrong_rate_raises(tmp_path):
    p = str(tmp_path / "c.wav")
    synth.write_wav_pcm16(p, np.zeros(100, np.int16), rate=16000)
    with pytest.raises(ValueError):
        wavio.read_wav(p)
    with pytest.raises(FileNotFoundError):
        wavio.read_wav(str(tmp_path / "missing.wav"))


def test_synth_is_seeded_and_has_trim_margin():
    a, b = synth.speech_like(1.3, 5), synth.speech_like(1.3, 5)
    assert np.array_equal(a, b) and a.dtype == np.float32 and a.size == int(round(1.3 * 22050))
    db = synth._trim_db(a)
    assert np.all(np.abs(db + 60.0) >= 3.0)
    d = synth.ljspeech_durations(13100, seed=1)
    assert d.min() >= 1.0 and d.max() <= 10.0 and 6.0 < d.mean() < 7.0


def test_npy_layout_written_like_prepare_data(tmp_path):
    """prepare_data.py:33-34 -> mel fortran_order True (payload [80][T]), mag C-order."""
    T = 7
    mel = np.arange(80 * T, dtype=np.float32).reshape(80, T).T
    mag = np.zeros((T, 1025), np.float32)
    np.save(str(tmp_path / "x.pt"), mel)
    np.save(str(tmp_path / "x.mag"), mag)
    hdr = open(str(tmp_path / "x.pt.npy"), "rb").read(128)
    assert b"'fortran_order': True" in hdr and b"'shape': (7, 80)" in hdr and b"'descr': '<f4'" in hdr
    hdr = open(str(tmp_path / "x.mag.npy"), "rb").read(128)
    assert b"'fortran_order': False" in hdr and b"'shape': (7, 1025)" in hdr


def _declared_symbols():
    names = set()
    inc = os.path.join(ROOT, "include")
    for fn in os.listdir(inc):
        if fn.endswith(".h"):
            src = open(os.path.join(inc, fn)).read()
            src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
            names |= set(re.findall(r"\b(unast_[a-z0-9_]+)\s*\(", src))
    return names


def test_c_abi_exports_every_declared_symbol():
    from unast_b200 import _lib
    declared = _declared_symbols()
    assert len(declared) >= 12
    assert declared == {s[0] for s in _lib.SYMBOLS}, "ctypes table and include/*.h disagree"
    assert os.path.exists(_lib.LIB_PATH), "build the library first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), "library does not export " + name
    assert lib.unast_abi_version() == 2                                               # no compute call without a GPU


def test_product_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from unast_b200 import utils as U
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        U.spectrograms_from_waves([np.zeros(4000, np.float32)])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        U.spectrogram2wav(np.zeros((8, 1025), np.float32))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "unast_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, fn)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), fn


_WORKER = r"""
import os, sys, json
import numpy as np
import torch.distributed as dist
sys.path.insert(0, %(root)r)
from unast_b200 import dist as D, synth
rank, local_rank, world = D.init_from_env(backend="gloo")
frames = np.round(synth.ljspeech_durations(400, seed=1) * 80.18).astype(int)
batches = D.my_batches(frames, rank, world, 64)
mine = [i for b in batches for i in b]
load = float(sum(frames[i] for i in mine))
total = D.sum_over_ranks(load)
slowest = D.max_over_ranks(load)
gathered = [None] * world
dist.all_gather_object(gathered, mine)
D.barrier()
if rank == 0:
    print(json.dumps({"total": total, "slowest": slowest, "owners": gathered, "expect": float(frames.sum())}))
dist.destroy_process_group()
"""


def test_world_size_2_sharding_over_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % {"root": ROOT})
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    line = [l for l in out.stdout.splitlines() if l.startswith("{")][-1]
    r = json.loads(line)
    assert r["total"] == r["expect"]
    owners = r["owners"]
    assert sorted(owners[0] + owners[1]) == list(range(400)) and not set(owners[0]) & set(owners[1])
    assert r["slowest"] / (r["total"] / 2) < 1.01                                   # balanced to < 1 %


def test_loader_sort_is_the_reference_permutation():
    """unast_b200.preprocess sorts like src/preprocess.py:98-104: descending text length, ties in original order."""
    from unast_b200 import preprocess as PP
    from oracle import spectral_oracle as so
    rng = np.random.RandomState(3)
    tlen = [5, 9, 5, 2, 9, 9, 1]
    batch = [{"text": rng.randint(1, 50, size=n).astype(np.int32), "text_length": n, "fname": str(i)} for i, n in enumerate(tlen)]
    order, tl = PP._sort_by_text_length(batch)
    ref = [i for i, _ in sorted(zip(range(len(tlen)), tlen), key=lambda x: x[1], reverse=True)]
    assert order == ref == [1, 4, 5, 0, 2, 3, 6] and tl == tlen
    padded = PP._prepare_data([batch[i]["text"] for i in order])
    assert np.array_equal(padded, so._prepare_data([batch[i]["text"] for i in order])) and padded.shape == (7, 9)


def test_long_form_chunks_cover_what_their_frames_see():
    """Frame-range sharding of one long waveform: ranges partition the frames; every kept frame's taps (and the sample
    before them, for the pre-emphasis) lie inside the chunk unless the chunk edge is the true signal edge."""
    from unast_b200 import longform as LF
    for n in (1101, 5000, 275 * 400 + 13, 13_230_000):
        T = 1 + n // 275
        for world in (1, 2, 3, 8):
            rs = LF.frame_ranges(T, world)
            assert rs[0][0] == 0 and rs[-1][1] == T and all(rs[i][1] == rs[i + 1][0] for i in range(world - 1))
            ra = LF.frame_ranges(T, world, align=LF.GL_BLOCK)                     # canonical Griffin-Lim blocks: aligned inner edges
            assert ra[0][0] == 0 and ra[-1][1] == T and all(ra[i][1] == ra[i + 1][0] for i in range(world - 1))
            assert all(e % LF.GL_BLOCK == 0 or e == T for _, e in ra[:-1]) and all(abs(x[1] - y[1]) <= LF.GL_BLOCK // 2 for x, y in zip(ra, rs))
            for a, b in rs:
                c0, c1, keep = LF.chunk_for_range(n, a, b)
                if b <= a:
                    continue
                assert c0 % 275 == 0 and keep.start == a - c0 // 275 and keep.stop - keep.start == b - a
                assert c0 == 0 or 275 * a - 552 >= c0
                assert c1 == n or 275 * (b - 1) + 551 <= c1 - 1
                assert 1 + (c1 - c0) // 275 >= keep.stop


def test_fragment_sample_ranges_tile_the_waveform():
    """unast_griffinlim_fragment_samples (host-only): the samples the ranks of the sharded long-form Griffin-Lim own are
    disjoint, ordered and cover [0, 275 (T - 1)) exactly, for every frame_ranges partition."""
    import ctypes
    from unast_b200 import _lib
    from unast_b200.longform import frame_ranges
    lib = ctypes.CDLL(_lib.LIB_PATH)
    fn = lib.unast_griffinlim_fragment_samples
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64)]
    for T in (48110, 4811, 73, 18):
        for world in (1, 2, 3, 8):
            if T // world < 9:
                continue
            pos = 0
            for a, b in frame_ranges(T, world):
                s, e = ctypes.c_int64(), ctypes.c_int64()
                assert fn(a, b - a, T, ctypes.byref(s), ctypes.byref(e)) == 0
                assert s.value == pos and e.value > s.value
                pos = e.value
            assert pos == 275 * (T - 1)
    s, e = ctypes.c_int64(), ctypes.c_int64()
    assert fn(5, 10, 12, ctypes.byref(s), ctypes.byref(e)) != 0                   # fragment beyond the utterance


def test_griffinlim_layout_is_consistent_with_workspace_bytes():
    """unast_griffinlim_layout (host-only): three rotating buffers of equal size inside the workspace, mailboxes before them."""
    import ctypes
    from unast_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    lib.unast_griffinlim_layout.restype = ctypes.c_int
    lib.unast_griffinlim_layout.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_void_p]
    lib.unast_griffinlim_workspace_bytes.restype = ctypes.c_size_t
    lib.unast_griffinlim_workspace_bytes.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_float]
    for frames, mom in (([24055], 0.0), ([100, 9, 4000], 0.99)):
        fr = np.asarray(frames, np.int32)
        out = (ctypes.c_size_t * 5)()
        assert lib.unast_griffinlim_layout(fr.ctypes.data_as(ctypes.c_void_p), fr.size, mom, ctypes.cast(out, ctypes.c_void_p)) == 0
        o0, o1, o2, mail, total = (int(v) for v in out)
        pad = int(sum(2048 + 275 * (n - 1) for n in frames)) * 4
        assert o1 - o0 == o2 - o1 >= pad and mail < o0 and o2 + (o1 - o0) <= total
        assert total == lib.unast_griffinlim_workspace_bytes(fr.ctypes.data_as(ctypes.c_void_p), fr.size, mom)
        assert o0 % 256 == 0 and mail % 256 == 0
