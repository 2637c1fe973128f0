"""CPU: the oracle (numpy restatement) against fixtures produced by the reference's own code
(tests/golden/make_golden.py ran /root/reference/src/utils.py and gl_vocoder.py over a torch-based
librosa stand-in) and against the reference's two shape known-answers."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, snr_db
from oracle import spectral_oracle as so

MEL_MAG_TOL = 1e-4      # north-star tolerance on the normalised spectrograms (values in [1e-8, 1])


@pytest.mark.parametrize("name", ["fwd_a", "fwd_b"])
def test_forward_matches_reference(golden, name):
    g = golden(name)
    y = g["pcm"].astype(np.float32) / np.float32(32768.0)                      # librosa.load of a PCM16 file
    mel, mag, bounds = so.spectrograms_from_wave(y, return_bounds=True)
    assert bounds == tuple(int(v) for v in g["bounds"])                          # discrete: exact
    assert mel.shape == g["mel"].shape and mag.shape == g["mag"].shape
    assert mel.dtype == np.float32 and mag.dtype == np.float32
    assert np.abs(mel - g["mel"]).max() <= 5e-6
    assert np.abs(mag - g["mag"]).max() <= 5e-6
    assert mel.shape[0] == 1 + (bounds[1] - bounds[0]) // so.hop_length         # T = 1 + N // 275


def test_memory_order_of_outputs(golden):
    """utils.py:275-276 + prepare_data.py:33-34: mel is F-contiguous (.npy fortran_order True), mag C-contiguous."""
    y = golden("fwd_b")["pcm"].astype(np.float32) / np.float32(32768.0)
    mel, mag = so.spectrograms_from_wave(y)
    assert mel.flags.f_contiguous and not mel.flags.c_contiguous
    assert mag.flags.c_contiguous


def test_classic_griffin_lim_matches_reference(golden):
    g = golden("gl_classic")
    wav = so.spectrogram2wav(g["mag_in"])                                        # ap.n_iter = 60, zero phase
    assert wav.dtype == np.float32 and wav.shape == g["wav"].shape
    assert snr_db(g["wav"], wav) >= 80.0


def test_fast_griffin_lim_matches_reference(golden):
    g = golden("gl_fast")
    np.random.seed(int(g["seed"]))                                                # librosa draws from the global RNG
    wav = so.vocode_one(g["mag_in"]).astype(np.float32)
    assert wav.shape == g["wav"].shape
    assert snr_db(g["wav"], wav) >= 80.0


def test_known_answer_lengths():
    # notebooks/testing_prepostnet.ipynb:593 -> 809 frames give 222,200 samples; samples/test.wav = 275*399
    assert so.invert_spectrogram(np.ones((1025, 809), np.complex64)).size == 222200
    assert so.invert_spectrogram(np.ones((1025, 400), np.complex64)).size == 109725
    y = np.zeros(220500, np.float32)
    y[::7] = 0.1
    assert so.stft(y).shape == (1025, 802)                                       # 10 s clip -> 802 frames


def test_summary_records_reference_agreement():
    with open(os.path.join(GOLDEN, "summary.json")) as fh:
        s = json.load(fh)["cases"]
    assert s["known_answer_809"]["istft_len_reference"] == 222200
    assert s["test_wav"]["known_answer_samples_275x399"] and s["test_wav"]["T"] == 400
    for k in ("fwd_a", "fwd_b", "test_wav"):
        assert s[k]["oracle_vs_reference_max_abs_mag"] <= 5e-6
        assert s[k]["oracle_vs_reference_max_abs_mel"] <= 5e-6
    assert s["gl_classic"]["oracle_vs_reference_snr_db"] >= 80
    assert s["gl_fast"]["oracle_vs_reference_snr_db"] >= 80


def test_mel_filterbank_properties():
    fb = so.mel_filterbank()
    assert fb.shape == (80, 1025) and fb.dtype == np.float32
    assert int((fb != 0).sum()) == 2001                                          # SURVEY A.4 probe
    assert np.all(fb >= 0) and np.all(fb.max(axis=1) > 0)
    assert np.all(fb[:, 1024] == 0) and np.all(fb[:, 0] == 0)
    assert abs(float(fb.max()) - 0.02415) < 1e-4


def test_trim_edge_cases():
    rng = np.random.RandomState(0)
    y = 1e-6 * rng.randn(6000).astype(np.float32)
    y[2000:3000] += 0.5 * rng.randn(1000).astype(np.float32)
    s, e = so.trim_bounds(y)
    assert s % 512 == 0 and 0 <= s <= 2000 and 3000 <= e <= 6000
    assert so.trim_bounds(np.zeros(4096, np.float32)) == (0, 4096)               # all frames equal the max -> all kept
    # numpy's reflect padding wraps repeatedly when the pad exceeds the signal (librosa 0.7.1 has no length check)
    assert so.trim_bounds(np.ones(1000, np.float32)) == (0, 1000)


def test_window_sumsquare_matches_direct_sum():
    T = 9
    wss = so.window_sumsquare(T)
    w2 = so.pad_center(so.hann_periodic() ** 2, 2048)
    direct = np.zeros(2048 + 275 * (T - 1))
    for t in range(T):
        direct[275 * t:275 * t + 2048] += w2
    assert np.abs(wss - direct).max() < 1e-6


def test_denormalise_formula():
    m = np.linspace(-0.2, 1.2, 50, dtype=np.float32)[None, :].repeat(3, 0)
    S = so.denormalise(m)
    ref = 10.0 ** (6.0 * np.clip(m.T.astype(np.float64), 0, 1) - 4.8)            # == (10^((m*100-80)/20))^1.2
    assert S.shape == (50, 3) and np.allclose(S, ref, rtol=2e-5)


def test_zero_phase_start_amplifies_last_bit_differences_of_S():
    """Why end-to-end Griffin-Lim parity is a statement about libm: numpy's SIMD float32 powf (what `denormalise` runs on
    this image, like the reference's own statements would) is not correctly rounded, and the zero-phase start of
    utils.griffin_lim (utils.py:311-313) turns last-bit differences of S into a visible difference of the waveform --
    the reference differs from ITSELF by this much across numpy / libm versions."""
    y = _speechy(22050, 11).astype(np.float32)
    mag = so.spectrograms_from_wave(y, do_trim=False)[1]
    S = so.denormalise(mag)                                                       # float32 statements of utils.py:289-298
    c = np.clip(mag.T, 0, 1)
    m = (c * np.float32(100) - np.float32(100) + np.float32(20)).astype(np.float32)
    a = np.power(10.0, (m * np.float32(0.05)).astype(np.float64)).astype(np.float32)    # correctly rounded 10^x
    S_cr = np.power(a.astype(np.float64), np.float64(np.float32(1.2))).astype(np.float32)   # correctly rounded a^1.2f
    ulp = np.abs(S.view(np.int32).astype(np.int64) - S_cr.view(np.int32).astype(np.int64))
    assert ulp.max() <= 4                                                          # the same numbers to the last few bits (measured: 3, 36 % of the bins)
    x0, x1 = so.istft(S.astype(np.complex64)), so.istft(S_cr.astype(np.complex64))
    d0 = 10 * np.log10((x0.astype(np.float64) ** 2).sum() / max(((x0.astype(np.float64) - x1) ** 2).sum(), 1e-300))
    if (ulp > 0).mean() > 0.01:                                                   # (a libm with correctly rounded powf: nothing to see)
        assert d0 < 135.0                                                          # float32 resolution would be ~ 150 dB
        g0, g1 = so.griffin_lim(S.astype(np.complex64), iters=30), so.griffin_lim(S_cr.astype(np.complex64), iters=30)
        d30 = 10 * np.log10((g0.astype(np.float64) ** 2).sum() / max(((g0.astype(np.float64) - g1) ** 2).sum(), 1e-300))
        assert d30 < 110.0


def test_short_clips_are_reflected_repeatedly():
    """numpy.pad(mode="reflect") under librosa.stft / effects.trim reflects a clip of <= 1024 samples more than once; the index
    rule the CUDA loader restates (period 2 (N - 1), N == 1: constant), checked against numpy itself, and the oracle runs on
    such clips (1 + N // 275 frames)."""
    def refl(m, N):
        if N == 1:
            return 0
        P = 2 * (N - 1)
        m %= P
        return m if m < N else P - m
    for N in (1, 2, 3, 275, 600, 1024, 1025):
        y = np.arange(N, dtype=np.float32) + 1
        ref = np.pad(y, 1024, mode="reflect")
        mine = np.array([y[refl(k - 1024, N)] for k in range(N + 2048)], np.float32)
        assert np.array_equal(ref, mine), N
    for N in (2, 300, 1024):
        w = (0.1 * np.random.RandomState(N).randn(N)).astype(np.float32)
        mel, mag, b = so.spectrograms_from_wave(w, return_bounds=True)
        assert mel.shape == (1 + (b[1] - b[0]) // 275, 80) and mag.shape[1] == 1025 and 0 <= b[0] < b[1] <= N


def test_deemphasis_inverts_preemphasis():
    rng = np.random.RandomState(1)
    y = rng.randn(5000)
    pre = np.append(y[0], y[1:] - 0.97 * y[:-1])
    assert np.abs(so.deemphasis(pre) - y).max() < 1e-9


def test_forward_chain_against_torchaudio_transforms():
    """Independent witness for the forward chain after the trim: torchaudio's own pre-emphasis, Spectrogram (its framing,
    reflect padding, zero-padded Hann), Slaney MelScale and amplitude_to_DB, in float64, against the oracle's float32 path --
    no code shared with oracle/ or with tests/golden/make_golden.py's stand-ins beyond torch.stft."""
    torch = pytest.importorskip("torch")
    ta = pytest.importorskip("torchaudio")
    y = _speechy(22050, 3).astype(np.float32)
    mel, mag = so.spectrograms_from_wave(y, do_trim=False)
    yt = torch.from_numpy(y.astype(np.float64))
    pre = torch.cat([yt[:1], ta.functional.preemphasis(yt, 0.97)[1:]])          # utils.py:251 keeps y[0]; torchaudio's first sample is y[0] too
    spec = ta.transforms.Spectrogram(n_fft=2048, win_length=1102, hop_length=275, power=1.0, center=True, pad_mode="reflect",
                                     window_fn=lambda n: torch.hann_window(n, periodic=True, dtype=torch.float64))(pre)
    fb = ta.functional.melscale_fbanks(1025, 0.0, 11025.0, 80, 22050, norm="slaney", mel_scale="slaney").to(torch.float64)
    mel_t = fb.T @ spec
    db = lambda x: ta.functional.amplitude_to_DB(x, 20.0, 1e-5, 0.0)             # 20 log10(max(1e-5, x))
    nrm = lambda d: torch.clamp((d - 20.0 + 100.0) / 100.0, 1e-8, 1.0)
    # float64 against the reference's float32 samples / complex64 spectra: the north-star gate (1e-4) on every bin -- the
    # largest differences sit in bins ~100 dB below the frame peak -- and rounding level on the typical one
    dg, dm = np.abs(nrm(db(spec)).numpy().T - mag), np.abs(nrm(db(mel_t)).numpy().T - mel)
    assert dg.max() <= 1e-4 and dm.max() <= 1e-4 and np.median(dg) <= 1e-6 and np.median(dm) <= 1e-6
    # de-emphasis (utils.py:301): torchaudio's IIR against scipy.signal.lfilter as the oracle calls it
    w = torch.from_numpy(np.random.RandomState(2).randn(4000))
    iir = ta.functional.lfilter(w, torch.tensor([1.0, -0.97], dtype=torch.float64), torch.tensor([1.0, 0.0], dtype=torch.float64), clamp=False)
    assert np.abs(iir.numpy() - so.deemphasis(w.numpy())).max() <= 1e-9     # (torchaudio's own deemphasis() clamps to [-1, 1])


# ---- next row: specaugment / noise_fn (src/utils.py:40-75), pinned on the reference's own output ----
def _seed(seed):
    import random
    import torch
    torch.manual_seed(int(seed))
    random.seed(int(seed))


def test_specaugment_restatement_matches_reference_output(golden):
    g = golden("augment")
    for key, rz in (("sa_mean", False), ("sa_zero", True)):
        _seed(g["seed"])
        spans = so.specaugment_spans(g["mel_len"])
        out = so.specaugment_apply(g["mel"], spans, replace_with_zero=rz)
        changed = out != g["mel"]
        assert np.array_equal(changed, g[key] != g["mel"])                      # the same rows are masked
        assert np.array_equal(out[~changed], g[key][~changed])
        if rz:
            assert np.array_equal(out, g[key])
        else:
            assert np.abs(out - g[key]).max() <= 6e-8                            # 1 ulp of a fill value near 0.5
    # both spans run along the time axis (whole rows), as `res[i][:][a:b]` does
    rows = (g["sa_zero"] != g["mel"]).any(axis=2)
    assert np.array_equal(rows[..., None] & (g["mel"] != 0), g["sa_zero"] != g["mel"])


def test_noise_fn_restatement_matches_reference_output(golden):
    g = golden("augment")
    assert np.array_equal(so.noise_apply(g["mel"], g["noise_mask"]), g["noise"])
    assert 0.2 < 1.0 - g["noise_mask"].mean() < 0.4                              # mask_p = .3


def test_specaugment_host_draws_and_error_match_reference(golden):
    from unast_b200.augment import draw_specaugment_spans
    g = golden("augment")
    _seed(g["seed"])
    want = so.specaugment_spans(g["mel_len"])
    _seed(g["seed"])
    assert np.array_equal(draw_specaugment_spans(g["mel_len"]), want)
    import torch
    _seed(g["seed"])
    assert np.array_equal(draw_specaugment_spans(torch.from_numpy(g["mel_len"])), want)
    _seed(g["seed"])
    with pytest.raises(ValueError) as e:
        draw_specaugment_spans(np.full(6, 10))
    assert "ValueError: %s" % e.value == str(g["short_error"])


# ---------------------------------------------------------------------------------------------------------------
# Independent witnesses (no code shared with oracle/): scipy / torch for the inverse transform, brute-force O(N T)
# restatements of librosa.filters.window_sumsquare and librosa.effects.trim straight from SURVEY.md Appendix A.
# ---------------------------------------------------------------------------------------------------------------
def _speechy(n, seed):
    rs = np.random.RandomState(seed)
    t = np.arange(n) / 22050.0
    y = sum(a * np.sin(2 * np.pi * f * t + p) for a, f, p in zip(rs.rand(12) * 0.1, 100 + 2000 * rs.rand(12), 6.28 * rs.rand(12)))
    return (y * (0.5 + 0.5 * np.sin(2 * np.pi * 3 * t)) + 1e-3 * rs.randn(n)).astype(np.float32)


def test_real_speech_fixture_matches_reference(golden):
    """samples/test.wav of the reference (float32 WAV, 275 * 399 samples) through its own get_spectrograms."""
    g = golden("real_test_wav")
    assert g["wave"].size == 275 * 399
    mel, mag, b = so.spectrograms_from_wave(g["wave"], return_bounds=True)
    assert b == tuple(int(v) for v in g["bounds"]) and mel.shape == g["mel"].shape
    assert np.abs(mel - g["mel"]).max() <= 5e-6
    assert np.abs(mag[g["mag_rows"]] - g["mag"]).max() <= 5e-6


def test_istft_against_torch_and_scipy():
    """oracle.istft vs torch.istft (float64) on every sample and vs scipy.signal.istft on interior samples."""
    import scipy.signal
    import torch
    y = _speechy(275 * 60, 5)
    X = so.stft(y)                                                               # [1025, 61] complex64
    mine = so.istft(X)
    assert mine.size == 275 * 60
    win = torch.zeros(2048, dtype=torch.float64)
    win[473:473 + 1102] = torch.from_numpy(0.5 - 0.5 * np.cos(2 * np.pi * np.arange(1102) / 1102))
    ref = torch.istft(torch.from_numpy(X.astype(np.complex128)), n_fft=2048, hop_length=275, win_length=2048, window=win,
                      center=True, length=275 * 60).numpy()
    assert np.abs(mine - ref).max() <= 2e-6
    # scipy: boundary=True trims nperseg // 2 like center=True; its scaling is window.sum() and it normalises by the
    # window-sum envelope as librosa does (interior samples only: scipy treats the edges differently)
    w = win.numpy()
    _, ys = scipy.signal.istft(X.astype(np.complex128) / w.sum(), fs=1.0, window=w, nperseg=2048, noverlap=2048 - 275, nfft=2048,
                               input_onesided=True, boundary=True)
    n = min(ys.size, mine.size)
    assert np.abs(mine[2048:n - 2048] - ys[2048:n - 2048]).max() <= 2e-6


def test_griffin_lim_loops_against_torchaudio():
    """Independent witness for the two Griffin-Lim loops: torchaudio.functional.griffinlim (its own stft / istft / momentum
    update, float64) started from the same phases -- momentum 0.99 == librosa.griffinlim as gl_vocoder.py:26 calls it,
    momentum 0 == the classic loop of utils.py:309-320 (the phase normalisations differ only below 1e-8)."""
    torch = pytest.importorskip("torch")
    taf = pytest.importorskip("torchaudio.functional")
    y = _speechy(275 * 80, 9)
    S = np.abs(so.stft(y)).astype(np.float32)                                    # [1025, 81]
    win = torch.zeros(2048, dtype=torch.float64)
    win[473:473 + 1102] = torch.from_numpy(0.5 - 0.5 * np.cos(2 * np.pi * np.arange(1102) / 1102))
    St = torch.from_numpy(S.astype(np.float64))

    def snr(a, b):
        a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
        return 10 * np.log10((a ** 2).sum() / max(((a - b) ** 2).sum(), 1e-300))

    for iters in (1, 4, 12):
        # rand_init=False starts torchaudio from unit phases: the zero-phase start of the reference / init=None of librosa
        ref_m = taf.griffinlim(St, win, 2048, 275, 2048, 1.0, iters, 0.99, 275 * 80, False).numpy()
        mine_m = so.librosa_griffinlim(S, iters=iters, momentum=0.99, angles0=np.ones(S.shape, np.complex64))
        assert mine_m.shape == ref_m.shape and snr(ref_m, mine_m) >= 85.0, (iters, snr(ref_m, mine_m))
        ref_c = taf.griffinlim(St, win, 2048, 275, 2048, 1.0, iters, 0.0, 275 * 80, False).numpy()
        mine_c = so.griffin_lim(S.astype(np.complex64), iters=iters)
        assert mine_c.shape == ref_c.shape and snr(ref_c, mine_c) >= 85.0, (iters, snr(ref_c, mine_c))


def test_window_sumsquare_brute_force():
    """Appendix A.3: wss[n] = sum_t hann^2[n - 275 t - 473] in float32 accumulation, by direct summation."""
    T = 23
    wss = so.window_sumsquare(T)
    h2 = (0.5 - 0.5 * np.cos(2 * np.pi * np.arange(1102) / 1102)) ** 2
    n = 2048 + 275 * (T - 1)
    assert wss.shape == (n,) and wss.dtype == np.float32
    brute = np.zeros(n, np.float64)
    for q in range(n):
        for t in range(T):
            i = q - 275 * t - 473
            if 0 <= i < 1102:
                brute[q] += h2[i]
    assert np.abs(wss - brute).max() <= 1e-5 * brute.max()
    assert np.all((wss > np.finfo(np.float32).tiny) == (brute > 0))              # same support: same samples are normalised


def test_trim_brute_force():
    """Appendix A.5 by direct loops (reflect index arithmetic written out, no np.pad / framing helper)."""
    rs = np.random.RandomState(9)
    y = np.concatenate([1e-5 * rs.randn(3000), 0.3 * rs.randn(9000), 1e-5 * rs.randn(2500)]).astype(np.float32)
    N = y.size
    nfr = 1 + N // 512
    mse = np.zeros(nfr)
    for f in range(nfr):
        acc = 0.0
        for j in range(2048):
            i = f * 512 + j - 1024
            if i < 0:
                i = -i
            if i >= N:
                i = 2 * (N - 1) - i
            acc += float(y[i]) ** 2
        mse[f] = acc / 2048
    db = 10 * np.log10(np.maximum(1e-10, mse)) - 10 * np.log10(max(1e-10, mse.max()))
    nz = np.flatnonzero(db > -60)
    expect = (int(nz[0]) * 512, min(N, (int(nz[-1]) + 1) * 512))
    assert so.trim_bounds(y) == expect
    assert 1000 < expect[0] <= 3072 and 12000 <= expect[1] < 14000              # the silences really were trimmed (frames reach +-1024)


def test_pcm16_fixture_of_vocode(golden):
    """gl_vocoder.py:34 sf.write -> PCM_16: the samples in the reference-written file vs the float data handed to it."""
    g = golden("gl_fast")
    assert g["pcm16"].dtype == np.int16 and g["pcm16"].shape == g["wav"].shape
    d = g["pcm16"].astype(np.float64) - g["wav"].astype(np.float64) * 32767.0
    d = (d + 32768.0) % 65536.0 - 32768.0                                        # clipping is off by default: |x| > 1 wraps
    assert np.abs(d).max() <= 0.51
