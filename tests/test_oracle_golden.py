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


def test_deemphasis_inverts_preemphasis():
    rng = np.random.RandomState(1)
    y = rng.randn(5000)
    pre = np.append(y[0], y[1:] - 0.97 * y[:-1])
    assert np.abs(so.deemphasis(pre) - y).max() < 1e-9


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
