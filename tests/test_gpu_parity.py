"""GPU (-m gpu): the CUDA path, called through the C ABI (ctypes -> libunast_spectral.so), against the
CPU oracle on identical seeded inputs.  Tolerances are the north star's:
  normalised mel / mag: max |d| <= 1e-4 (values in [1e-8, 1]); trim bounds, T and lengths exact;
  Griffin-Lim: waveform SNR >= 40 dB vs the oracle's output from the same initial phases, spectral
  convergence within 1e-3 of the oracle's."""
import os

import numpy as np
import pytest

from conftest import snr_db
from oracle import spectral_oracle as so

pytestmark = pytest.mark.gpu

TOL = 1e-4
GL_SNR_DB = 40.0
GL_SC_TOL = 1e-3


def _torch():
    import torch
    return torch


def _clips(specs):
    from unast_b200 import synth
    return [synth.speech_like(d, s) for d, s in specs]


# ------------------------------------------------------------------------------------------------
# FFT core on its own
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("prec,tol", [("f64", 2e-6), ("f32", 5e-4)])
def test_rfft_core_matches_numpy(engine, prec, tol):
    torch = _torch()
    rng = np.random.RandomState(0)
    F = 37                                                     # ragged vs the 4-frames-per-CTA launch
    fr = np.zeros((F, 2048), np.float32)
    fr[:, :1102] = rng.randn(F, 1102).astype(np.float32)
    ref = np.fft.rfft(fr.astype(np.float64), axis=1)
    out = engine.debug_rfft(torch.from_numpy(fr).cuda(), prec).cpu().numpy()
    assert np.abs(out - ref).max() <= tol * np.abs(ref).max()
    # linearity (size-independent property): rfft(a + 2 b) = rfft(a) + 2 rfft(b)
    a, b = fr[:16], fr[16:32]
    lhs = engine.debug_rfft(torch.from_numpy(a + 2 * b).cuda(), prec).cpu().numpy()
    rhs = out[:16] + 2 * out[16:32]
    assert np.abs(lhs - rhs).max() <= 4 * tol * np.abs(ref).max()


def test_irfft_core_matches_numpy(engine):
    torch = _torch()
    rng = np.random.RandomState(1)
    fr = np.zeros((9, 2048))
    fr[:, :1102] = rng.randn(9, 1102)
    spec = np.fft.rfft(fr, axis=1).astype(np.complex64)
    back = engine.debug_irfft(torch.from_numpy(spec).cuda()).cpu().numpy()
    ref = np.fft.irfft(spec.astype(np.complex128), axis=1)[:, :1102]
    assert np.abs(back - ref).max() <= 2e-6 * max(1.0, np.abs(ref).max())


def test_mel_basis_equals_oracle(engine):
    assert np.abs(engine.mel_basis() - so.mel_filterbank()).max() <= 1e-9


# ------------------------------------------------------------------------------------------------
# forward path
# ------------------------------------------------------------------------------------------------
def test_frontend_f64_parity_ragged_batch():
    from unast_b200 import utils as U
    waves = _clips([(10.0, 0), (3.3, 1), (1.0, 2), (6.57, 3), (2.2, 4)])
    res = U.spectrograms_from_waves(waves, precision="f64", return_bounds=True)
    assert len(res) == len(waves)
    for w, (mel, mag, b) in zip(waves, res):
        omel, omag, ob = so.spectrograms_from_wave(w, return_bounds=True)
        assert b == ob                                                           # discrete: exact
        assert mel.shape == omel.shape and mag.shape == omag.shape
        assert mel.dtype == np.float32 and mag.dtype == np.float32
        assert mel.flags.f_contiguous and mag.flags.c_contiguous               # utils.py:275-276 memory order
        assert np.abs(mag - omag).max() <= TOL
        assert np.abs(mel - omel).max() <= TOL


def test_frontend_f32_fast_mode_error_is_bounded():
    """The float32 FFT mode is NOT the parity mode: documented bound (DESIGN.md 'precision')."""
    from unast_b200 import utils as U
    waves = _clips([(4.0, 20), (2.0, 21)])
    for w, (mel, mag) in zip(waves, U.spectrograms_from_waves(waves, precision="f32")):
        omel, omag = so.spectrograms_from_wave(w)
        d = np.abs(mag - omag)
        assert d.max() <= 2e-3 and np.quantile(d, 0.999) <= 5e-5
        assert np.abs(mel - omel).max() <= TOL                                   # mel sums ~25 bins: well inside


def test_frontend_matches_reference_golden(golden):
    from unast_b200 import utils as U
    for name in ("fwd_a", "fwd_b"):
        g = golden(name)
        (mel, mag, b), = U.spectrograms_from_waves([g["pcm"]], return_bounds=True)   # int16 input path
        assert b == tuple(int(v) for v in g["bounds"])
        assert mel.shape == g["mel"].shape
        assert np.abs(mel - g["mel"]).max() <= TOL and np.abs(mag - g["mag"]).max() <= TOL


def test_frontend_int16_equals_float_input():
    from unast_b200 import synth, utils as U
    w = _clips([(2.5, 30)])[0]
    q = synth.to_pcm16(w)
    (mel_i, mag_i), = U.spectrograms_from_waves([q])
    (mel_f, mag_f), = U.spectrograms_from_waves([q.astype(np.float32) / np.float32(32768.0)])
    assert np.array_equal(mel_i, mel_f) and np.array_equal(mag_i, mag_f)


def test_frontend_int16_prefetch_alignments():
    """PCM16 batch whose utterances start at every alignment inside a 16-byte chunk (the cp.async sample prefetch of the
    float64 kernel copies aligned chunks and picks the samples out by offset): bit-identical to the float input path."""
    from unast_b200 import synth, utils as U
    rng = np.random.RandomState(11)
    lens = [275 * 14 + 3 + k for k in range(9)] + [275 * 30 + 1, 275 * 40 + 6]
    qs = [synth.to_pcm16(0.3 * rng.randn(n).astype(np.float32)) for n in lens]
    res_i = U.spectrograms_from_waves(qs, do_trim=False)
    res_f = U.spectrograms_from_waves([q.astype(np.float32) / np.float32(32768.0) for q in qs], do_trim=False)
    for (mel_i, mag_i), (mel_f, mag_f) in zip(res_i, res_f):
        assert np.array_equal(mel_i, mel_f) and np.array_equal(mag_i, mag_f)


def test_frontend_without_trim_and_minimum_length():
    from unast_b200 import utils as U
    rng = np.random.RandomState(5)
    waves = [0.1 * rng.randn(n).astype(np.float32) for n in (1025, 1375, 2048, 5000)]
    res = U.spectrograms_from_waves(waves, do_trim=False)
    for w, (mel, mag) in zip(waves, res):
        omel, omag = so.spectrograms_from_wave(w, do_trim=False)
        assert mel.shape == omel.shape == (1 + w.size // 275, 80)
        assert np.abs(mag - omag).max() <= TOL and np.abs(mel - omel).max() <= TOL


def test_frontend_edge_cases_raise():
    from unast_b200 import utils as U
    assert U.spectrograms_from_waves([]) == []
    with pytest.raises(IndexError):
        U.spectrograms_from_waves([np.zeros(0, np.float32)])                     # utils.py:251 raises on empty audio


def test_frontend_short_clips_are_reflected_like_numpy_pad():
    """Clips of <= 1024 samples: numpy.pad(mode="reflect") reflects them more than once inside librosa's 1024-sample padding
    (stft and effects.trim alike); the drop-in does the same instead of refusing them."""
    from unast_b200 import synth, utils as U
    rng = np.random.RandomState(9)
    waves = [0.1 * rng.randn(n).astype(np.float32) for n in (2, 3, 275, 300, 600, 1023, 1024)] + [np.ones(600, np.float32)]
    for do_trim in (False, True):
        res = U.spectrograms_from_waves(waves, do_trim=do_trim, return_bounds=True)
        for w, (mel, mag, b) in zip(waves, res):
            omel, omag, ob = so.spectrograms_from_wave(w, do_trim=do_trim, return_bounds=True)
            assert b == ob and mel.shape == omel.shape and mag.shape == omag.shape, (w.size, b, ob)
            assert np.abs(mag - omag).max() <= TOL and np.abs(mel - omel).max() <= TOL, w.size
    pcm = [synth.to_pcm16(w) for w in waves[2:7]]                                  # the int16 path too
    for w, (mel, mag) in zip(pcm, U.spectrograms_from_waves(pcm, do_trim=False)):
        omel, omag = so.spectrograms_from_wave(w.astype(np.float32) / np.float32(32768.0), do_trim=False)
        assert np.abs(mag - omag).max() <= TOL and np.abs(mel - omel).max() <= TOL


def test_frontend_padded_loader_layout(engine):
    """[B, T_max, 80] zero-padded batch tensor (collate_fn_transformer, preprocess.py:82-121)."""
    torch = _torch()
    waves = _clips([(1.5, 40), (2.5, 41), (1.1, 42)])
    lens = np.array([w.size for w in waves])
    off = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)).cuda()
    wave = torch.from_numpy(np.concatenate(waves)).cuda()
    bounds = engine.trim_bounds(wave, off, int(lens.max()))
    fo = engine.frame_offsets(bounds)
    engine.check()
    fo_h = fo.cpu().numpy()
    T_max = int(np.diff(fo_h).max())
    mel, mag = engine.frontend(wave, off, bounds, fo, int(fo_h[-1]), precision="f64", want_mag=False,
                               mel_layout="padded", T_max=T_max)
    assert mag is None and tuple(mel.shape) == (3, T_max, 80)
    mel = mel.cpu().numpy()
    for b, w in enumerate(waves):
        omel, _ = so.spectrograms_from_wave(w)
        T = omel.shape[0]
        assert np.abs(mel[b, :T] - omel).max() <= TOL
        assert np.all(mel[b, T:] == 0)


def test_get_spectrograms_drop_in_and_npy_files(tmp_path):
    from unast_b200 import synth, utils as U
    from unast_b200.prepare_data import PrepareDataset
    ids = ["LJ001-0001", "LJ001-0002", "LJ001-0003"]
    waves = _clips([(1.2, 50), (2.0, 51), (1.6, 52)])
    for i, w in zip(ids, waves):
        synth.write_wav_pcm16(str(tmp_path / (i + ".wav")), w)
    (tmp_path / "metadata.csv").write_text("".join("%s|text|text\n" % i for i in ids))
    mel, mag = U.get_spectrograms(str(tmp_path / (ids[0] + ".wav")))             # reference signature
    q = synth.to_pcm16(waves[0]).astype(np.float32) / np.float32(32768.0)
    omel, omag = so.spectrograms_from_wave(q)
    assert mel.shape == omel.shape and np.abs(mel - omel).max() <= TOL and np.abs(mag - omag).max() <= TOL
    with pytest.raises(FileNotFoundError):
        U.get_spectrograms(str(tmp_path / "nope.wav"))
    ds = PrepareDataset(str(tmp_path / "metadata.csv"), str(tmp_path))
    assert len(ds) == 3
    ds.process([0, 1, 2])
    for i, w in zip(ids, waves):
        hdr = open(str(tmp_path / (i + ".pt.npy")), "rb").read(128)
        assert b"'fortran_order': True" in hdr                                    # same bytes as prepare_data.py:33
        m, g = np.load(str(tmp_path / (i + ".pt.npy"))), np.load(str(tmp_path / (i + ".mag.npy")))
        q = synth.to_pcm16(w).astype(np.float32) / np.float32(32768.0)
        omel, omag = so.spectrograms_from_wave(q)
        assert m.shape == omel.shape and g.shape == omag.shape
        assert np.abs(m - omel).max() <= TOL and np.abs(g - omag).max() <= TOL


def test_frontend_full_size_batch_properties():
    """256 LJSpeech-shaped clips (C2 sample): frame counts exact for every clip, values in range,
    a seeded subset checked against the oracle, and scale invariance: x -> x/2 lowers every unclipped
    normalised value by exactly 20 log10(2) / 100."""
    from unast_b200 import synth, utils as U
    durs = synth.ljspeech_durations(256, seed=1)
    waves = [synth.speech_like(float(d), 1000 + i) for i, d in enumerate(durs)]
    res = U.spectrograms_from_waves(waves, return_bounds=True)
    for w, (mel, mag, b) in zip(waves, res):
        assert mel.shape[0] == mag.shape[0] == 1 + (b[1] - b[0]) // 275
        assert mag.min() >= 1e-8 and mag.max() <= 1.0 and np.isfinite(mel).all()
    for i in np.random.RandomState(0).choice(256, 6, replace=False):
        omel, omag, ob = so.spectrograms_from_wave(waves[i], return_bounds=True)
        assert res[i][2] == ob
        assert np.abs(res[i][0] - omel).max() <= TOL and np.abs(res[i][1] - omag).max() <= TOL
    half = U.spectrograms_from_waves([w * np.float32(0.5) for w in waves[:8]])
    for (mel, mag, _), (mel2, mag2) in zip(res[:8], half):
        inside = (mag > 0.1) & (mag < 0.99)
        assert np.abs((mag - mag2)[inside] - 0.2 * np.log10(2.0)).max() <= 2e-6


# ------------------------------------------------------------------------------------------------
# inverse path
# ------------------------------------------------------------------------------------------------
def _mags(specs):
    return [so.spectrograms_from_wave(w)[1] for w in _clips(specs)]


def _check_gl(y, ref, S):
    assert y.shape == ref.shape
    assert snr_db(ref, y) >= GL_SNR_DB
    assert abs(so.spectral_convergence(y, S) - so.spectral_convergence(ref, S)) <= GL_SC_TOL


@pytest.mark.parametrize("iters", [0, 1, 5, 30])
def test_griffin_lim_classic_zero_phase(iters):
    from unast_b200 import utils as U
    mags = _mags([(2.0, 10), (1.2, 11), (3.1, 12)])
    outs = U.spectrogram2wav_batch(mags, n_iter=iters, do_trim=False)
    for m, y in zip(mags, outs):
        ref = so.spectrogram2wav(m, iters=iters, return_untrimmed=True)
        assert y.dtype == np.float32 and y.size == 275 * (m.shape[0] - 1)
        _check_gl(y, ref, so.denormalise(m))


def test_griffin_lim_classic_100_iterations_seeded_phase():
    from unast_b200 import utils as U
    mags = _mags([(1.5, 13), (1.0, 14)])
    rs = np.random.RandomState(3)
    a0 = [np.exp(2j * np.pi * rs.rand(1025, m.shape[0])) for m in mags]
    outs = U.spectrogram2wav_batch(mags, n_iter=100, angles0=[a.T for a in a0], do_trim=False)
    for m, a, y in zip(mags, a0, outs):
        ref = so.spectrogram2wav(m, iters=100, angles0=a, return_untrimmed=True)
        _check_gl(y, ref, so.denormalise(m))


def test_griffin_lim_momentum_seeded_phase():
    """librosa.griffinlim variant of gl_vocoder.py:26 (32 iterations, momentum 0.99) from shared seeded phases."""
    from unast_b200.gl_vocoder import vocode_arrays
    mags = _mags([(2.0, 10), (1.2, 11)])
    rs = np.random.RandomState(3)
    a0 = [np.exp(2j * np.pi * rs.rand(1025, m.shape[0])) for m in mags]
    outs = vocode_arrays(mags, n_iter=32, momentum=0.99, angles0=[a.T for a in a0], do_trim=False)
    for m, a, y in zip(mags, a0, outs):
        ref = so.vocode_one(m, iters=32, momentum=0.99, angles0=a, return_untrimmed=True).astype(np.float32)
        _check_gl(y, ref, so.denormalise(m))


def test_griffin_lim_matches_reference_golden(golden):
    from unast_b200 import utils as U
    from unast_b200.gl_vocoder import vocode_arrays
    g = golden("gl_classic")
    y = U.spectrogram2wav(g["mag_in"])                                           # reference signature; ap.n_iter = 60
    assert y.dtype == np.float32 and y.shape == g["wav"].shape                   # trim bounds equal
    assert snr_db(g["wav"], y) >= GL_SNR_DB
    g = golden("gl_fast")
    np.random.seed(int(g["seed"]))                                               # global RNG, as librosa draws it
    y, = vocode_arrays([g["mag_in"]])
    assert y.shape == g["wav"].shape and snr_db(g["wav"], y) >= GL_SNR_DB


def test_griffin_lim_trim_and_minimum_frames():
    from unast_b200 import utils as U
    mags = _mags([(1.6, 15)])
    mags.append(np.ascontiguousarray(mags[0][:5]))                               # T = 5 is the smallest legal spectrogram
    outs = U.spectrogram2wav_batch(mags, n_iter=8)
    for m, y in zip(mags, outs):
        ref = so.spectrogram2wav(m, iters=8)
        assert y.shape == ref.shape                                              # librosa.effects.trim bounds agree
        assert snr_db(ref, y) >= GL_SNR_DB
    with pytest.raises(ValueError):
        U.spectrogram2wav_batch([mags[0][:4]], n_iter=1)
    with pytest.raises(ValueError):
        U.spectrogram2wav_batch([np.zeros((8, 80), np.float32)], n_iter=1)


def test_invert_spectrogram_and_griffin_lim_raw():
    """utils.invert_spectrogram (:323-328) and utils.griffin_lim (:309-320) called directly."""
    from unast_b200 import utils as U
    w = _clips([(1.3, 16)])[0]
    X = so.stft(w)
    y = U.invert_spectrogram(X)
    ref = so.invert_spectrogram(X)
    assert y.shape == ref.shape and snr_db(ref, y) >= 100.0
    S = np.abs(X).astype(np.float32)
    y = U.griffin_lim(S)                                                          # 60 iterations, zero phase
    ref = so.griffin_lim(S)
    _check_gl(y, np.asarray(ref, np.float32), S)


def test_stft_istft_round_trip_full_length():
    """Size-independent property at full size: istft(stft(x)) == x away from the edges (COLA holds for
    hann 1102 / hop 275 after window-sum normalisation), on a 10 s clip."""
    from unast_b200 import utils as U
    w = _clips([(10.0, 17)])[0]
    X = so.stft(w)                                                                # complex STFT from the oracle ...
    y = U.invert_spectrogram(X)                                                   # ... inverted on the GPU
    n = y.size
    assert n == 275 * (X.shape[1] - 1)
    assert np.abs(y[2048:n - 2048] - w[2048:n - 2048]).max() <= 2e-5


def test_vocode_cli_files(tmp_path):
    from unast_b200 import gl_vocoder, wavio
    mags = _mags([(1.2, 18), (1.4, 19)])
    (tmp_path / "m").mkdir()
    for i, m in enumerate(mags):
        np.save(str(tmp_path / "m" / ("utt%d.mag.npy" % i)), m)
    (tmp_path / "list.csv").write_text("utt0|a|a\nutt1|b|b\n")
    np.random.seed(11)
    gl_vocoder.vocode(str(tmp_path / "list.csv"), str(tmp_path / "m"), str(tmp_path / "out"))
    np.random.seed(11)
    for i, m in enumerate(mags):
        y, rate = wavio.read_wav(str(tmp_path / "out" / ("utt%d.wav" % i)))
        ref = so.vocode_one(m).astype(np.float32)                                 # draws rand(1025, T) in list order
        assert y.dtype == np.int16                                                # sf.write default subtype: PCM_16
        y = y.astype(np.float32) / np.float32(32767.0)
        assert rate == 22050 and y.shape == ref.shape and snr_db(ref, y) >= GL_SNR_DB


# ------------------------------------------------------------------------------------------------
# structure-specific edge cases of the kernels (tile / run boundaries, short utterances)
# ------------------------------------------------------------------------------------------------
def test_frontend_tile_boundaries_ragged_frame_counts():
    """Frame counts around multiples of the 12-frame (float64 kernel) and 16-frame (float32 kernel) tiles and a batch
    large enough that several CTAs share an utterance: every frame must land in the right row."""
    from unast_b200 import utils as U
    rng = np.random.RandomState(7)
    lens = [275 * (t - 1) + int(rng.randint(0, 275)) for t in (11, 12, 13, 15, 16, 17, 23, 24, 25, 31, 32, 33, 5, 48)]
    waves = [0.2 * rng.randn(n).astype(np.float32) for n in lens]
    res = U.spectrograms_from_waves(waves, do_trim=False)
    for w, (mel, mag) in zip(waves, res):
        omel, omag = so.spectrograms_from_wave(w, do_trim=False)
        assert mel.shape == omel.shape and mag.shape == omag.shape
        assert np.abs(mag - omag).max() <= TOL and np.abs(mel - omel).max() <= TOL


def test_griffin_lim_run_seams_and_short_utterances():
    """Griffin-Lim on a batch that mixes very short spectrograms (5..13 frames: exact edge envelope, runs shorter
    than the 827-sample overlap) with a long one (every warp's run has seams on both sides)."""
    from unast_b200 import utils as U
    long_mag = _mags([(6.0, 60)])[0]
    mags = [np.ascontiguousarray(long_mag[40:40 + t]) for t in (5, 6, 7, 8, 11, 12, 13)] + [long_mag]
    outs = U.spectrogram2wav_batch(mags, n_iter=12, do_trim=False)
    for m, y in zip(mags, outs):
        ref = so.spectrogram2wav(m, iters=12, return_untrimmed=True)
        _check_gl(y, ref, so.denormalise(m))
    # the same long spectrogram alone (small grid: runs of ~8 frames) gives the same waveform as inside the batch
    alone = U.spectrogram2wav_batch([long_mag], n_iter=12, do_trim=False)[0]
    assert snr_db(outs[-1], alone) >= 80.0


def test_griffin_lim_is_deterministic():
    from unast_b200 import utils as U
    mags = _mags([(3.0, 61), (1.0, 62)])
    a = U.spectrogram2wav_batch(mags, n_iter=10, do_trim=False)
    b = U.spectrogram2wav_batch(mags, n_iter=10, do_trim=False)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_loader_collate_on_the_fly_matches_reference_collate():
    """collate_fn_transformer (src/preprocess.py:82-121) with mels computed on the fly: order (text length descending,
    ties stable), zero padding, dtypes and values against the oracle's line-by-line restatement."""
    torch = _torch()
    from unast_b200 import preprocess as PP
    rng = np.random.RandomState(21)
    waves = _clips([(1.3, 60), (2.2, 61), (1.0, 62), (1.8, 63), (1.5, 64)])
    tlen = [7, 12, 7, 3, 12]                                              # ties: stable order inside equal lengths
    texts = [rng.randint(1, 70, size=n).astype(np.int32) for n in tlen]
    names = ["a", "b", "c", "d", "e"]
    ref, ref_names = so.collate_fn_transformer(so.loader_items_from_waves(texts, waves, names))
    batch = [{"text": tx, "wave": w, "text_length": len(tx), "fname": n} for tx, w, n in zip(texts, waves, names)]
    (text, mel, text_length, mel_length), got_names = PP.collate_fn_transformer(batch)
    assert got_names == ref_names == ["b", "e", "a", "c", "d"]
    assert text.dtype == torch.long and mel.dtype == torch.float32 and mel.is_cuda
    assert np.array_equal(text.cpu().numpy(), ref[0]) and np.array_equal(text_length.cpu().numpy(), ref[2])
    assert np.array_equal(mel_length.cpu().numpy(), ref[3])
    assert tuple(mel.shape) == ref[1].shape
    m = mel.cpu().numpy()
    assert np.abs(m - ref[1]).max() <= TOL
    for b, T in enumerate(ref[3]):
        assert np.all(m[b, T:] == 0)
    # mel-only branch of collate_fn_postnet: batch order kept
    mel2, lens2, names2 = PP.collate_fn_postnet_mel(batch)
    assert names2 == names and lens2 == [so.spectrograms_from_wave(w)[0].shape[0] for w in waves]
    assert np.abs(mel2.cpu().numpy()[0, :lens2[0]] - so.spectrograms_from_wave(waves[0])[0]).max() <= TOL


def test_long_form_frame_range_sharding_is_bit_identical():
    """C5: one long waveform, frames sharded over 'ranks' with a halo (unast_b200.longform): the concatenation of the
    ranks' frames is bit-identical to the single-GPU result and within the parity gate of the oracle."""
    from unast_b200 import longform as LF, synth, utils as U
    w = np.concatenate([synth.speech_like(9.0, 70 + i) for i in range(4)])          # 36 s
    (mel1, mag1), = U.spectrograms_from_waves([w])
    for world in (1, 3, 5):
        parts = [LF.spectrograms_long(w, rank=r, world=world) for r in range(world)]
        T = parts[0][3]
        assert [p[2] for p in parts] == LF.frame_ranges(T, world) and T == mel1.shape[0]
        mel = np.concatenate([p[0] for p in parts], axis=0)
        mag = np.concatenate([p[1] for p in parts], axis=0)
        assert np.array_equal(mel, mel1) and np.array_equal(mag, mag1)
    omel, omag = so.spectrograms_from_wave(w)
    assert np.abs(mel1 - omel).max() <= TOL and np.abs(mag1 - omag).max() <= TOL


# ------------------------------------------------------------------------------------------------
# next row: specaugment / noise_fn on the padded mel batch
# ------------------------------------------------------------------------------------------------
def _seed_all(seed):
    import random
    torch = _torch()
    torch.manual_seed(int(seed))
    random.seed(int(seed))


def test_specaugment_matches_reference_output(golden):
    """Same seeds as the fixture: the drop-in must mask the reference's rows with the reference's fill value."""
    torch = _torch()
    from unast_b200.augment import specaugment
    g = golden("augment")
    mel = torch.from_numpy(g["mel"]).cuda()
    keep = mel.clone()
    for key, rz in (("sa_mean", False), ("sa_zero", True)):
        _seed_all(g["seed"])
        out = specaugment(mel, torch.from_numpy(g["mel_len"]), replace_with_zero=rz)
        assert out.data_ptr() != mel.data_ptr() and torch.equal(mel, keep)        # a clone; the input is untouched
        out = out.cpu().numpy()
        assert np.array_equal(out != g["mel"], g[key] != g["mel"])
        assert np.abs(out - g[key]).max() <= (0.0 if rz else 6e-8)
    _seed_all(g["seed"])
    with pytest.raises(ValueError):                                              # utils.py:66 on a too-short utterance
        specaugment(mel[:, :10].contiguous(), torch.full((6,), 10, dtype=torch.int64))


def test_noise_fn_matches_reference_formula_on_the_device_stream():
    torch = _torch()
    from unast_b200.augment import noise_fn
    rng = np.random.RandomState(11)
    for shape in ((5, 77, 80), (3, 50, 257), (2, 9, 3)):                        # float4 path and the scalar path
        x = torch.from_numpy(rng.randn(*shape).astype(np.float32)).cuda()
        torch.manual_seed(99)
        out = noise_fn(x, mask_p=.3)
        torch.manual_seed(99)                                                    # utils.py:46-49, same generator
        gen = torch.zeros(shape[:2], device=x.device).fill_(1 - .3)
        want = x * torch.bernoulli(gen).unsqueeze(-1)
        assert torch.equal(out, want)
    dropped = (out.abs().sum(-1) == 0).float().mean().item()
    assert 0.0 <= dropped <= 1.0


def test_mel_augment_large_batch_against_oracle(engine):
    """Ragged batch at loader scale, spans + keep mask in one pass, against the oracle's row arithmetic."""
    torch = _torch()
    rng = np.random.RandomState(5)
    B, T, D = 37, 811, 80
    lens = rng.randint(300, T + 1, size=B)
    lens[0] = T
    x = rng.uniform(0.0, 1.0, size=(B, T, D)).astype(np.float32)
    for i, n in enumerate(lens):
        x[i, n:] = 0.0
    _seed_all(3)
    spans = so.specaugment_spans(lens)
    keep = (rng.rand(B, T) > 0.3).astype(np.float32)
    want = so.noise_apply(so.specaugment_apply(x, spans), keep)
    xd = torch.from_numpy(x).cuda()
    got = engine.mel_augment(xd, spans=torch.from_numpy(spans).cuda(), keep=torch.from_numpy(keep).cuda()).cpu().numpy()
    assert np.abs(got - want).max() <= 6e-8
    assert np.array_equal(got == 0, want == 0)
    # determinism of the two-stage mean, and in-place operation where it is allowed
    again = engine.mel_augment(xd, spans=torch.from_numpy(spans).cuda(), keep=torch.from_numpy(keep).cuda()).cpu().numpy()
    assert np.array_equal(got, again)
    z = xd.clone()
    engine.mel_augment(z, spans=torch.from_numpy(spans).cuda(), replace_with_zero=True, out=z)
    assert np.array_equal(z.cpu().numpy(), so.specaugment_apply(x, spans, replace_with_zero=True))
    with pytest.raises(Exception):
        engine.mel_augment(xd, spans=torch.from_numpy(spans).cuda(), out=xd)   # the mean is read from x


def test_loader_batch_then_specaugment_like_the_supervised_step():
    """collate_fn_transformer -> specaugment(mel, mel_len) (train.py:232-236) without leaving the device."""
    torch = _torch()
    from unast_b200 import preprocess, synth
    from unast_b200.augment import specaugment
    items = [{"text": np.arange(3 + i), "wave": synth.speech_like(1.6 + 0.2 * i, 60 + i)} for i in range(4)]
    text, mel, text_len, mel_len = preprocess.collate_fn_transformer(items)
    _seed_all(8)
    aug = specaugment(mel, mel_len)
    _seed_all(8)
    spans = so.specaugment_spans(mel_len.cpu().numpy())
    want = so.specaugment_apply(mel.cpu().numpy(), spans)
    assert np.abs(aug.cpu().numpy() - want).max() <= 6e-8


# ------------------------------------------------------------------------------------------------
# round 2: real speech, benchmark-scale Griffin-Lim, long waveforms, gradients, device handling
# ------------------------------------------------------------------------------------------------
def test_real_speech_fixture_through_the_cuda_path(golden):
    """The reference's only real-speech file (Transformer-TTS/samples/test.wav, float32 input path) against the mel /
    mag its own get_spectrograms produced (tests/golden/make_golden.py)."""
    from unast_b200 import utils as U
    g = golden("real_test_wav")
    (mel, mag, b), = U.spectrograms_from_waves([g["wave"]], return_bounds=True)
    assert b == tuple(int(v) for v in g["bounds"]) and mel.shape == g["mel"].shape
    assert np.abs(mel - g["mel"]).max() <= TOL
    assert np.abs(mag[g["mag_rows"]] - g["mag"]).max() <= TOL
    omel, omag = so.spectrograms_from_wave(g["wave"])                             # and every bin against the oracle
    assert np.abs(mag - omag).max() <= TOL and np.abs(mel - omel).max() <= TOL


def test_griffin_lim_benchmark_scale_batch():
    """C3 as benchmarked: 64 spectrograms of LJSpeech-shaped lengths in ONE call, 100 classic iterations from zero
    phase; four of them (shortest, longest, first, last) against the oracle."""
    from unast_b200 import synth, utils as U
    durs = synth.ljspeech_durations(64, seed=2)
    waves = [synth.speech_like(float(d), 3000 + i) for i, d in enumerate(durs)]
    (mags) = [m for _, m in U.spectrograms_from_waves(waves)]
    outs = U.spectrogram2wav_batch(mags, n_iter=100, do_trim=False)
    pick = sorted({0, 63, int(np.argmin(durs)), int(np.argmax(durs))})
    for i in pick:
        ref = so.spectrogram2wav(mags[i], iters=100, return_untrimmed=True)
        _check_gl(outs[i], ref, so.denormalise(mags[i]))


def test_griffin_lim_from_the_oracles_target_magnitudes():
    """The kernel alone: both sides start from the SAME target magnitudes S (the oracle's, passed through with power 0).  The
    zero-phase start w * irfft(S) is computed in float64 (gl_init_f64_kernel): it is the tail of a large cancellation, and
    the one step of the whole run whose float32 rounding decides the distance to the reference.  End to end (the CUDA path
    de-normalising the magnitudes itself) the last-bit differences between numpy's SIMD float32 powf and correctly rounded
    powers enter that same step: the tests above hold 40 dB there, this one holds 65 dB (measured 73-112 dB) and 130 dB at
    zero iterations."""
    from unast_b200 import utils as U
    mags = _mags([(1.5, 1021), (2.7, 1005), (6.0, 1042)])
    Ss = [np.ascontiguousarray(so.denormalise(m).T) for m in mags]
    for iters, gate in ((0, 130.0), (100, 65.0)):
        outs = U._gl_batch(Ss, iters, 0.0, None, 0.0, True, False)
        for m, y in zip(mags, outs):
            ref = so.spectrogram2wav(m, iters=iters, return_untrimmed=True)
            assert y.shape == ref.shape and snr_db(ref, y) >= gate, (iters, snr_db(ref, y))


def test_griffin_lim_long_waveform():
    """One 60 s spectrogram (4,811 frames: hundreds of warp runs and seams inside one utterance), 30 iterations."""
    from unast_b200 import synth, utils as U
    w = np.concatenate([synth.speech_like(10.0, 80 + i) for i in range(6)])
    (mel, mag), = U.spectrograms_from_waves([w], do_trim=False)
    y, = U.spectrogram2wav_batch([mag], n_iter=30, do_trim=False)
    ref = so.spectrogram2wav(mag, iters=30, return_untrimmed=True)
    _check_gl(y, ref, so.denormalise(mag))


def test_noise_fn_lets_gradients_through():
    """utils.py:49 is `to_noise * zero_mask`: differentiable, and applied to the embedded phonemes (network.py:431)."""
    torch = _torch()
    from unast_b200.augment import noise_fn
    x = torch.randn(4, 33, 16, device="cuda", requires_grad=True)
    torch.manual_seed(5)
    y = noise_fn(x, mask_p=.3)
    torch.manual_seed(5)
    keep = torch.bernoulli(torch.zeros(4, 33, device="cuda").fill_(.7)).unsqueeze(-1)
    w = torch.randn_like(x)
    (y * w).sum().backward()
    assert torch.equal(y.detach(), x.detach() * keep) and torch.equal(x.grad, w * keep)
    with pytest.raises(TypeError):
        noise_fn(x.double())


def test_unaligned_pcm16_view_is_rejected(engine):
    """The PCM16 path prefetches 16-byte chunks relative to the base pointer: a sliced int16 tensor must be refused
    (UNAST_E_ARG), not fault."""
    torch = _torch()
    from unast_b200 import synth
    w = torch.from_numpy(synth.to_pcm16(synth.speech_like(1.5, 90))).cuda()
    v = w[3:]
    off = torch.tensor([0, v.numel()], dtype=torch.int64, device="cuda")
    bounds = torch.tensor([[0, v.numel()]], dtype=torch.int32, device="cuda")
    fo = engine.frame_offsets(bounds)
    with pytest.raises(Exception):
        engine.frontend(v, off, bounds, fo, int(fo[-1]))


def test_engine_on_a_non_current_device():
    """unast_create / every entry point select the context's device and restore the caller's."""
    torch = _torch()
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from unast_b200.engine import SpectralEngine
    from unast_b200 import synth
    assert torch.cuda.current_device() == 0
    eng = SpectralEngine(1)
    assert torch.cuda.current_device() == 0
    w = synth.speech_like(1.5, 91)
    wave = torch.from_numpy(w).to("cuda:1")
    off = torch.tensor([0, w.size], dtype=torch.int64, device="cuda:1")
    with torch.cuda.device(1):
        bounds = eng.trim_bounds(wave, off, w.size)
        fo = eng.frame_offsets(bounds)
        eng.check()
        mel, mag = eng.frontend(wave, off, bounds, fo, int(fo.cpu()[-1]))
    assert torch.cuda.current_device() == 0
    omel, omag = so.spectrograms_from_wave(w)
    assert np.abs(mag.cpu().numpy() - omag).max() <= TOL


def test_sharded_griffin_lim_over_two_gpus():
    """C5 inverse for real: two processes, one GPU each, seams linked over NVLink peer mappings (tools/longform_gl_canonical_mp.py
    under torchrun).  Canonical blocks: the gathered waveform must equal the single-GPU run bit for bit at 0 / 1 / 2 / 10 / 100
    iterations, twice in a row (short calls are the ones that exposed a missing hand-over after the last iteration)."""
    import subprocess
    import sys
    torch = _torch()
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(root, "tools", "longform_gl_canonical_mp.py"), "60"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=root)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [l for l in res.stdout.splitlines() if l.startswith("iters ")]
    assert len(lines) == 5, res.stdout[-2000:]
    for l in lines:
        assert "run-to-run equal True" in l and "mismatches 0 " in l, l


def test_sharded_griffin_lim_protocol_emulated_on_one_gpu():
    """C5 inverse: one long spectrogram cut into fragments with linked seams (what the ranks of the sharded run do over
    NVLink), emulated in one launch.  One fragment == the ordinary path bit for bit; 2 / 3 / 5 fragments differ from it
    only by the summation order at the run boundaries."""
    torch = _torch()
    from unast_b200 import longform as LF, synth, utils as U
    w = np.concatenate([synth.speech_like(8.0, 120 + i) for i in range(3)])            # 24 s
    (mel, mag), = U.spectrograms_from_waves([w], do_trim=False)
    n_iter = 12
    base, = U._gl_batch([mag], n_iter, 0.0, None, 1.2, False, False)                  # normalised, no de-emphasis / trim
    mag_d = torch.from_numpy(mag).cuda()
    one = LF.griffinlim_fragments_emulated(mag_d, 1, n_iter).cpu().numpy()
    assert np.array_equal(one, base)
    for world in (2, 3, 5):
        y = LF.griffinlim_fragments_emulated(mag_d, world, n_iter).cpu().numpy()
        assert y.shape == base.shape and snr_db(base, y) >= 80.0
    rs = np.random.RandomState(4)                                                      # momentum variant, seeded phases
    a0 = np.exp(2j * np.pi * rs.rand(mag.shape[0], 1025)).astype(np.complex64)
    a0_d = torch.from_numpy(a0).cuda()
    ref = LF.griffinlim_fragments_emulated(mag_d, 1, 8, momentum=0.99, angles0=a0_d).cpu().numpy()
    y = LF.griffinlim_fragments_emulated(mag_d, 4, 8, momentum=0.99, angles0=a0_d).cpu().numpy()
    assert snr_db(ref, y) >= 80.0
    # epilogue of the sharded run: de-emphasis + trim of the gathered waveform == the ordinary call's
    full, = U.spectrogram2wav_batch([mag], n_iter=n_iter, do_trim=True)
    eng = U.get_engine()
    L = base.size
    wav, bounds = eng.deemph_trim(torch.from_numpy(base).cuda(), torch.tensor([0, L], dtype=torch.int64, device="cuda"), L)
    s, e = (int(v) for v in bounds.cpu().numpy()[0])
    assert np.array_equal(wav.cpu().numpy()[s:e], full)


def test_sharded_griffin_lim_canonical_blocks_are_bit_identical_for_every_cut():
    """unast_griffinlim_set_block: with canonical summation blocks the overlap-add no longer depends on how the frames are cut
    into runs, fragments and ranks -- 1 / 2 / 3 / 5 / 7 fragments give the same bits (classic and momentum), the ordinary
    whole-utterance entry point too, and the result stays within rounding of the speed-optimised cut."""
    torch = _torch()
    from unast_b200 import longform as LF, synth, utils as U
    w = np.concatenate([synth.speech_like(8.0, 140 + i) for i in range(3)])            # 24 s, 1,925 frames (not a multiple of 4)
    (mel, mag), = U.spectrograms_from_waves([w], do_trim=False)
    n_iter = 15
    mag_d = torch.from_numpy(mag).cuda()
    one = LF.griffinlim_fragments_emulated(mag_d, 1, n_iter, canonical=True).cpu().numpy()
    for world in (2, 3, 5, 7):
        y = LF.griffinlim_fragments_emulated(mag_d, world, n_iter, canonical=True).cpu().numpy()
        assert np.array_equal(y, one), world
    eng = U.get_engine()
    eng.set_gl_block(LF.GL_BLOCK)
    try:
        whole, = U._gl_batch([mag], n_iter, 0.0, None, 1.2, False, False)            # ordinary entry point, canonical blocks
        pair = U._gl_batch([mag[:700], mag], n_iter, 0.0, None, 1.2, False, False)   # another batch composition: other run cuts
    finally:
        eng.set_gl_block(0)
    assert np.array_equal(whole, one) and np.array_equal(pair[1], one)
    fast, = U._gl_batch([mag], n_iter, 0.0, None, 1.2, False, False)
    assert snr_db(fast, one) >= 80.0
    rs = np.random.RandomState(4)                                                      # momentum variant, seeded phases
    a0_d = torch.from_numpy(np.exp(2j * np.pi * rs.rand(mag.shape[0], 1025)).astype(np.complex64)).cuda()
    ref = LF.griffinlim_fragments_emulated(mag_d, 1, 8, momentum=0.99, angles0=a0_d, canonical=True).cpu().numpy()
    for world in (2, 4):
        y = LF.griffinlim_fragments_emulated(mag_d, world, 8, momentum=0.99, angles0=a0_d, canonical=True).cpu().numpy()
        assert np.array_equal(y, ref), world


def test_prepare_streams_files_like_np_save(tmp_path):
    """prepare_data.prepare(): reader thread -> pinned staging -> GPU -> writer threads; the .npy files must be byte-identical
    to what np.save writes for the arrays get_spectrograms returns (prepare_data.py:33-34), over several chunks."""
    from unast_b200 import prepare_data, synth, utils as U
    ids = ["LJ%03d" % i for i in range(11)]
    waves = _clips([(1.1 + 0.13 * i, 200 + i) for i in range(11)])
    for i, w in zip(ids, waves):
        synth.write_wav_pcm16(str(tmp_path / (i + ".wav")), w)
    (tmp_path / "metadata.csv").write_text("".join("%s|x|x\n" % i for i in ids))
    n = prepare_data.prepare(str(tmp_path / "metadata.csv"), str(tmp_path), batch_size=3, writers=2)   # 4 chunks in flight
    assert n == 11
    for i in ids:
        mel, mag = U.get_spectrograms(str(tmp_path / (i + ".wav")))
        np.save(str(tmp_path / "ref.pt"), mel)
        np.save(str(tmp_path / "ref.mag"), mag)
        assert open(str(tmp_path / (i + ".pt.npy")), "rb").read() == open(str(tmp_path / "ref.pt.npy"), "rb").read()
        assert open(str(tmp_path / (i + ".mag.npy")), "rb").read() == open(str(tmp_path / "ref.mag.npy"), "rb").read()


def test_vocode_tensors_device_entry_matches_file_path():
    """Device-tensor hand-off (inf_vocoder.py:48-64 -> gl_vocoder.py:15-32 without the .npy round trip): a padded
    [B, T_max, 1025] device batch gives the waveforms vocode_arrays gives for the same magnitudes and seed."""
    torch = _torch()
    from unast_b200 import gl_vocoder as GV
    mags = _mags([(1.3, 31), (1.9, 32), (1.1, 33)])
    frames = [m.shape[0] for m in mags]
    pad = np.zeros((3, max(frames), 1025), np.float32)
    for b, m in enumerate(mags):
        pad[b, :frames[b]] = m
    np.random.seed(5)
    ref = GV.vocode_arrays(mags)
    np.random.seed(5)
    wav, off, bounds = GV.vocode_tensors(torch.from_numpy(pad).cuda(), frames)
    wav, bounds = wav.cpu().numpy(), bounds.cpu().numpy()
    for b in range(3):
        y = wav[off[b]:off[b + 1]][bounds[b, 0]:bounds[b, 1]]
        assert np.array_equal(y, ref[b])


def test_prenet_fc1_on_tensor_cores_matches_float32_torch(engine):
    """masking -> Linear(80 -> 256 / 128) + bias -> ReLU (module.py:93-95) in one tcgen05 kernel against plain float32 torch:
    operand splitting keeps float32 fidelity (2e-6 of the row's |x| . |w| scale; torch's own float32 GEMM is within 4x)."""
    torch = _torch()
    torch.backends.cuda.matmul.allow_tf32 = False
    rng = np.random.RandomState(12)
    for (B, T, N) in ((5, 333, 256), (3, 129, 128), (1, 7, 256), (2, 300, 96), (2, 70, 32), (41, 1003, 256)):   # the last: 3 tiles per CTA
        x = torch.from_numpy(rng.uniform(0, 1, size=(B, T, 80)).astype(np.float32)).cuda()
        W = torch.from_numpy((rng.randn(N, 80) / 9).astype(np.float32)).cuda()
        bias = torch.from_numpy(rng.randn(N).astype(np.float32)).cuda()
        lens = np.full(B, T)
        _seed_all(21)
        spans = torch.from_numpy(so.specaugment_spans(lens) if T > 120 else np.zeros((B, 4), np.int32)).cuda()
        keep = (torch.rand((B, T), device="cuda") > 0.3).float()
        packed, n_out = engine.prenet_pack(W, bias)
        y = engine.mel_prenet_fc1(x, packed, n_out, spans=spans, keep=keep)
        xm = engine.mel_augment(x, spans=spans, keep=keep)                          # the masking kernel, already pinned
        ref = torch.relu(torch.nn.functional.linear(xm.double(), W.double(), bias.double())).float()
        scale = (xm.abs() @ W.abs().T + bias.abs()).clamp_min(1e-6)
        assert y.shape == ref.shape
        assert ((y - ref).abs() / scale).max().item() <= 2e-6
        pre = engine.mel_prenet_fc1(x, packed, n_out, relu=False)                    # no masks, pre-activation
        ref2 = torch.nn.functional.linear(x.double(), W.double(), bias.double()).float()
        assert ((pre - ref2).abs() / (x.abs() @ W.abs().T + bias.abs())).max().item() <= 2e-6
        ref32 = torch.nn.functional.linear(x, W, bias)                               # torch's own float32 GEMM is no closer
        # float32-class, not TF32-class (a single TF32 pass is ~1e-3): what remains is the tensor cores' truncating float32
        # accumulation over 30 instructions (~20 ulp of the sum), within 10x of cuBLAS's float32 GEMM
        assert (pre - ref2).abs().max().item() <= 10 * (ref32 - ref2).abs().max().item() + 1e-6


def test_guard_bands_around_every_output(engine):
    """compute-sanitizer is closed on this pool (profiles/r3_sanitizer.txt): every kernel output is placed inside a larger
    canary-filled buffer and the canaries on both sides must survive -- at ragged sizes that straddle tile and run ends."""
    torch = _torch()
    from unast_b200 import synth
    CAN = 12345.0
    G = 4096

    def banded(n):
        buf = torch.full((n + 2 * G,), CAN, dtype=torch.float32, device="cuda")
        return buf, buf[G:G + n]

    def intact(buf, n):
        return bool((buf[:G] == CAN).all().item()) and bool((buf[G + n:] == CAN).all().item())

    waves = [synth.to_pcm16(synth.speech_like(d, 400 + i)) for i, d in enumerate((1.01, 2.37, 1.5, 3.03))]
    lens = np.array([w.size for w in waves])
    off = torch.from_numpy(np.concatenate([[0], np.cumsum(lens)])).cuda()
    wave = torch.from_numpy(np.concatenate(waves)).cuda()
    bounds = engine.trim_bounds(wave, off, int(lens.max()))
    fo = engine.frame_offsets(bounds)
    engine.check()
    F = int(fo.cpu()[-1])
    for prec in ("f64", "f32"):
        bm, mel = banded(80 * F)
        bg, mag = banded(F * 1025)
        engine.frontend(wave, off, bounds, fo, F, precision=prec, mel_out=mel, mag_out=mag.view(F, 1025))
        torch.cuda.synchronize()
        assert intact(bm, 80 * F) and intact(bg, F * 1025)
        assert float(mel.max()) <= 1.0 and float(mag.max()) <= 1.0                     # every element was written
    frames = np.diff(fo.cpu().numpy()).astype(np.int32)
    nw = int((275 * (frames.astype(np.int64) - 1)).sum())
    for mom in (0.0, 0.99):
        bw, wav = banded(nw)
        engine.griffinlim(mag.view(F, 1025).contiguous(), frames, n_iter=3, momentum=mom, wav_out=wav)
        torch.cuda.synchronize()
        assert intact(bw, nw) and bool(torch.isfinite(wav).all().item()) and float(wav.abs().max()) < 100.0
    B, T = 3, 131
    x = torch.rand((B, T, 80), device="cuda")
    spans = torch.tensor([[3, 5, 60, 40], [0, 0, 10, 99], [120, 11, 0, 1]], dtype=torch.int32, device="cuda")
    keep = (torch.rand((B, T), device="cuda") > 0.3).float()
    bo, out = banded(B * T * 80)
    engine.mel_augment(x, spans=spans, keep=keep, out=out.view(B, T, 80))
    torch.cuda.synchronize()
    assert intact(bo, B * T * 80)
    packed, n_out = engine.prenet_pack(torch.randn(256, 80, device="cuda"), torch.randn(256, device="cuda"))
    bp, pre = banded(B * T * 256)
    engine.mel_prenet_fc1(x, packed, n_out, spans=spans, keep=keep, out=pre.view(B, T, 256))
    torch.cuda.synchronize()
    assert intact(bp, B * T * 256) and float(pre.max()) < CAN
