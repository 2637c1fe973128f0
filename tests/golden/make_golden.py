"""Generates tests/golden/*.npz by running the REFERENCE'S OWN code.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py

What runs: /root/reference/src/utils.py (get_spectrograms :235-278, spectrogram2wav :281-306,
griffin_lim :309-320, invert_spectrogram :323-328) and /root/reference/src/gl_vocoder.py (vocode
:10-34), imported unmodified from where they lie.  Their third-party dependencies are not installed
in this image (librosa, soundfile, jiwer, eng_to_ipa), so this script registers
stand-in modules before the import:

  librosa    -- load / effects.trim / stft / istft / filters.mel / griffinlim rebuilt on
                torch.stft / torch.istft (float64) and torchaudio.functional.melscale_fbanks
                (norm='slaney', mel_scale='slaney').  This shares no code with oracle/ (which
                restates librosa on numpy.fft), so the fixtures are a second, independent witness
                of the librosa 0.7.1 semantics written out in SURVEY.md Appendix A.
  soundfile  -- write() as PCM_16 (soundfile's default subtype for .wav; libsndfile rounding lrint(x * 0x7FFF)) via
                scipy.io.wavfile.
  jiwer, data -- inert stubs (only imported at module top, never called on this path).

Outputs (committed, small):
  fwd_<name>.npz   pcm (int16 input), mel [T,80], mag [T,1025], bounds of the trim
  real_test_wav.npz  the reference's only real-speech file (Transformer-TTS/samples/test.wav, float32 WAV): wave, reference
                   mel [T,80], every 4th row of the reference mag, trim bounds
  gl_classic.npz   mag_in [T,1025], wav = reference spectrogram2wav (60 iterations, zero phase)
  gl_fast.npz      mag_in, seed, wav = reference gl_vocoder.vocode output (librosa.griffinlim defaults,
                   global np.random seeded with `seed` immediately before the call); pcm16 = the samples that land in the file
  summary.json     shapes, known-answer checks, oracle-vs-reference deviations measured at generation time
"""
from __future__ import annotations

import json
import os
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_SRC = "/root/reference/src"
sys.path.insert(0, ROOT)


# ---------------------------------------------------------------------------------------------
# librosa stand-in (torch / torchaudio based)
# ---------------------------------------------------------------------------------------------
def _make_librosa():
    import scipy.io.wavfile as wavfile
    import torchaudio

    lib = types.ModuleType("librosa")
    lib.__version__ = "0.7.1-standin"

    def load(path, sr=22050):
        rate, data = wavfile.read(path)
        if rate != sr:
            raise ValueError("resampling not supported by the stand-in")
        if data.dtype == np.int16:
            y = data.astype(np.float32) / np.float32(32768.0)
        elif data.dtype == np.int32:
            y = (data.astype(np.float64) / 2147483648.0).astype(np.float32)
        else:
            y = data.astype(np.float32)
        if y.ndim > 1:
            y = y.mean(axis=1).astype(np.float32)
        return y, rate

    def _hann(win_length):
        return torch.hann_window(win_length, periodic=True, dtype=torch.float64)

    def stft(y, n_fft=2048, hop_length=None, win_length=None, window="hann", center=True,
             dtype=np.complex64, pad_mode="reflect"):
        win_length = win_length or n_fft
        hop_length = hop_length or win_length // 4
        t = torch.from_numpy(np.ascontiguousarray(y)).to(torch.float64)
        X = torch.stft(t, n_fft, hop_length=hop_length, win_length=win_length, window=_hann(win_length),
                       center=center, pad_mode=pad_mode, normalized=False, onesided=True, return_complex=True)
        return np.asfortranarray(X.numpy().astype(dtype))

    def _wss(n_frames, hop_length, win_length, n_fft, dtype):
        w = _hann(win_length).numpy() ** 2
        lpad = (n_fft - win_length) // 2
        wp = np.zeros(n_fft)
        wp[lpad:lpad + win_length] = w
        n = n_fft + hop_length * (n_frames - 1)
        x = np.zeros(n, dtype=dtype)
        for i in range(n_frames):
            x[i * hop_length:i * hop_length + n_fft] += wp
        return x

    def istft(stft_matrix, hop_length=None, win_length=None, window="hann", center=True, dtype=np.float32,
              length=None):
        n_fft = 2 * (stft_matrix.shape[0] - 1)
        win_length = win_length or n_fft
        hop_length = hop_length or win_length // 4
        # frame synthesis through torch's inverse real FFT (float64); overlap-add and envelope
        # normalisation follow librosa (float32 accumulator), SURVEY.md Appendix A.3
        X = torch.from_numpy(np.ascontiguousarray(stft_matrix).astype(np.complex128))
        frames = torch.fft.irfft(X, n=n_fft, dim=0).numpy()
        lpad = (n_fft - win_length) // 2
        w = np.zeros(n_fft)
        w[lpad:lpad + win_length] = _hann(win_length).numpy()
        T = stft_matrix.shape[1]
        y = np.zeros(n_fft + hop_length * (T - 1), dtype=dtype)
        for t in range(T):
            y[t * hop_length:t * hop_length + n_fft] += w * frames[:, t]
        wss = _wss(T, hop_length, win_length, n_fft, dtype)
        nz = wss > np.finfo(dtype).tiny
        y[nz] /= wss[nz]
        return y[n_fft // 2:-(n_fft // 2)]

    def griffinlim(S, n_iter=32, hop_length=None, win_length=None, window="hann", center=True, dtype=np.float32,
                   length=None, pad_mode="reflect", momentum=0.99, init="random", random_state=None):
        rng = np.random if random_state is None else np.random.RandomState(seed=random_state)
        n_fft = 2 * (S.shape[0] - 1)
        angles = np.empty(S.shape, dtype=np.complex64)
        angles[:] = np.exp(2j * np.pi * rng.rand(*S.shape))
        rebuilt = 0.0
        for _ in range(n_iter):
            tprev = rebuilt
            inverse = istft(S * angles, hop_length=hop_length, win_length=win_length, dtype=dtype)
            rebuilt = stft(inverse, n_fft=n_fft, hop_length=hop_length, win_length=win_length)
            angles[:] = rebuilt - (momentum / (1 + momentum)) * tprev
            angles[:] /= np.abs(angles) + 1e-16
        return istft(S * angles, hop_length=hop_length, win_length=win_length, dtype=dtype)

    effects = types.ModuleType("librosa.effects")

    def trim(y, top_db=60, ref=np.max, frame_length=2048, hop_length=512):
        t = torch.from_numpy(np.ascontiguousarray(y)).to(torch.float64)
        tp = torch.nn.functional.pad(t[None, None], (frame_length // 2, frame_length // 2), mode="reflect")[0, 0]
        fr = tp.unfold(0, frame_length, hop_length)                       # (n_frames, frame_length)
        mse = (fr * fr).mean(dim=1).numpy()
        db = 10.0 * np.log10(np.maximum(1e-10, mse))
        db = db - 10.0 * np.log10(np.maximum(1e-10, ref(mse)))
        nz = np.flatnonzero(db > -top_db)
        if nz.size:
            start, end = int(nz[0]) * hop_length, min(y.shape[-1], (int(nz[-1]) + 1) * hop_length)
        else:
            start, end = 0, 0
        return y[start:end], np.asarray([start, end])

    effects.trim = trim
    filters = types.ModuleType("librosa.filters")

    def mel(sr, n_fft, n_mels=128):
        fb = torchaudio.functional.melscale_fbanks(n_freqs=1 + n_fft // 2, f_min=0.0, f_max=sr / 2.0, n_mels=n_mels,
                                                   sample_rate=sr, norm="slaney", mel_scale="slaney")
        return fb.T.contiguous().numpy().astype(np.float32)

    filters.mel = mel
    lib.load, lib.stft, lib.istft, lib.griffinlim = load, stft, istft, griffinlim
    lib.effects, lib.filters = effects, filters
    return lib, effects, filters


SF_LAST = {}


def _install_stubs():
    import scipy.io.wavfile as wavfile
    lib, effects, filters = _make_librosa()
    sys.modules["librosa"], sys.modules["librosa.effects"], sys.modules["librosa.filters"] = lib, effects, filters
    sf = types.ModuleType("soundfile")

    def sf_write(path, data, rate):
        # python-soundfile: default subtype for .wav is PCM_16 whatever the dtype; libsndfile (normalised floats, clipping
        # off) stores lrint(x * 0x7FFF) truncated to 16 bits.  The float data is kept so that the fixture can hold both.
        d = np.asarray(data, np.float64)
        SF_LAST["data"] = d.copy()
        wavfile.write(path, rate, (np.rint(d * 32767.0).astype(np.int64) & 0xFFFF).astype(np.uint16).view(np.int16))
    sf.write = sf_write
    sys.modules["soundfile"] = sf
    jw = types.ModuleType("jiwer")
    jw.wer = lambda *a, **k: 0.0
    sys.modules["jiwer"] = jw
    data = types.ModuleType("data")
    data.sequence_to_text = lambda seq: ""
    sys.modules["data"] = data


def _snr_db(ref, x):
    ref = np.asarray(ref, np.float64)
    x = np.asarray(x, np.float64)
    n = min(ref.size, x.size)
    err = ref[:n] - x[:n]
    return float(10 * np.log10((ref[:n] ** 2).sum() / max((err ** 2).sum(), 1e-300)))


def main():
    _install_stubs()
    sys.path.insert(0, REF_SRC)
    import utils as ref_utils            # /root/reference/src/utils.py, unmodified
    import gl_vocoder as ref_glv         # /root/reference/src/gl_vocoder.py, unmodified
    import audio_parameters as ref_ap

    from oracle import spectral_oracle as so
    from unast_b200 import synth

    summary = {"reference_commit": "7a5abb5", "reference_files": ["src/utils.py:235-328", "src/gl_vocoder.py:10-34"],
               "standin": "torch %s / torchaudio stft, istft, melscale_fbanks" % torch.__version__, "cases": {}}
    tmp = tempfile.mkdtemp(prefix="unast_golden_")

    # ---- forward path: synthetic clips written as PCM16 wav, read by the reference ----
    fwd = {"a": (1.7, 101), "b": (1.05, 102)}
    mags = {}
    for name, (dur, seed) in fwd.items():
        y = synth.speech_like(dur, seed)
        pcm = synth.to_pcm16(y)
        path = os.path.join(tmp, name + ".wav")
        synth.write_wav_pcm16(path, pcm)
        mel, mag = ref_utils.get_spectrograms(path)
        yf = pcm.astype(np.float32) / np.float32(32768.0)
        _, b = sys.modules["librosa"].effects.trim(yf)
        omel, omag, ob = so.spectrograms_from_wave(yf, return_bounds=True)
        assert mel.shape == omel.shape and tuple(int(v) for v in b) == ob, (mel.shape, omel.shape, b, ob)
        assert mel.flags.f_contiguous and mag.flags.c_contiguous               # SURVEY 8(b) memory order
        np.savez_compressed(os.path.join(HERE, "fwd_%s.npz" % name), pcm=pcm, mel=np.ascontiguousarray(mel),
                            mag=mag, bounds=np.asarray(b, np.int32))
        mags[name] = mag
        summary["cases"]["fwd_" + name] = {
            "samples": int(pcm.size), "bounds": [int(v) for v in b], "T": int(mel.shape[0]),
            "oracle_vs_reference_max_abs_mel": float(np.abs(mel - omel).max()),
            "oracle_vs_reference_max_abs_mag": float(np.abs(mag - omag).max())}

    # ---- real speech: samples/test.wav (not copied into the repo; deviations recorded only) ----
    tw = "/root/reference/src/Transformer-TTS/samples/test.wav"
    mel, mag = ref_utils.get_spectrograms(tw)
    omel, omag = so.get_spectrograms(tw)
    y_tw, _ = so.load(tw)
    _, b_tw = sys.modules["librosa"].effects.trim(y_tw)
    rows = np.arange(0, mag.shape[0], 4)                                  # every 4th magnitude row keeps the fixture small
    np.savez_compressed(os.path.join(HERE, "real_test_wav.npz"), wave=np.asarray(y_tw, np.float32), mel=np.ascontiguousarray(mel),
                        mag_rows=rows.astype(np.int32), mag=np.ascontiguousarray(mag[rows]), bounds=np.asarray(b_tw, np.int32))
    summary["cases"]["test_wav"] = {
        "samples": int(y_tw.size), "known_answer_samples_275x399": bool(y_tw.size == 275 * 399),
        "T": int(mel.shape[0]), "oracle_T": int(omel.shape[0]),
        "oracle_vs_reference_max_abs_mel": float(np.abs(mel - omel).max()) if mel.shape == omel.shape else None,
        "oracle_vs_reference_max_abs_mag": float(np.abs(mag - omag).max()) if mag.shape == omag.shape else None}

    # ---- inverse path, utils.spectrogram2wav (classic GL, ap.n_iter = 60, zero phase) ----
    mag_in = np.ascontiguousarray(mags["a"][:48])
    assert ref_ap.n_iter == 60
    wav = ref_utils.spectrogram2wav(mag_in)
    owav = so.spectrogram2wav(mag_in)
    S = so.denormalise(mag_in)
    np.savez_compressed(os.path.join(HERE, "gl_classic.npz"), mag_in=mag_in, wav=wav)
    summary["cases"]["gl_classic"] = {
        "T": 48, "n_iter": 60, "len": int(wav.size), "oracle_len": int(owav.size),
        "oracle_vs_reference_snr_db": _snr_db(wav, owav) if wav.size == owav.size else None,
        "spectral_convergence_reference": so.spectral_convergence(ref_utils.griffin_lim(S), S),
        "spectral_convergence_oracle": so.spectral_convergence(so.griffin_lim(S), S)}

    # known-answer: 809 frames -> 222,200 samples before trimming (notebooks/testing_prepostnet.ipynb:593)
    X = np.ones((1025, 809), np.complex64)
    summary["cases"]["known_answer_809"] = {"istft_len_reference": int(ref_utils.invert_spectrogram(X).size),
                                            "istft_len_oracle": int(so.invert_spectrogram(X).size), "expected": 222200}

    # ---- inverse path, gl_vocoder.vocode (librosa.griffinlim defaults; global RNG seeded here) ----
    seed = 1234
    mdir, odir = os.path.join(tmp, "mags"), os.path.join(tmp, "wavs")
    os.makedirs(mdir)
    mag_fast = np.ascontiguousarray(mags["b"][:40])
    np.save(os.path.join(mdir, "utt0.mag.npy"), mag_fast)
    with open(os.path.join(tmp, "list.csv"), "w") as fh:
        fh.write("utt0|some text|some text\n")
    np.random.seed(seed)
    ref_glv.vocode(os.path.join(tmp, "list.csv"), mdir, odir)
    import scipy.io.wavfile as wavfile
    rate, pcm_fast = wavfile.read(os.path.join(odir, "utt0.wav"))
    assert pcm_fast.dtype == np.int16
    wav_fast = SF_LAST["data"]                                          # the float64 array gl_vocoder.py:34 hands to sf.write
    np.random.seed(seed)
    owav_fast = so.vocode_one(mag_fast)
    np.savez_compressed(os.path.join(HERE, "gl_fast.npz"), mag_in=mag_fast, seed=np.int64(seed),
                        wav=wav_fast.astype(np.float32), pcm16=pcm_fast)
    summary["cases"]["gl_fast"] = {
        "T": 40, "n_iter": 32, "momentum": 0.99, "rate": int(rate), "len": int(wav_fast.size), "oracle_len": int(owav_fast.size),
        "oracle_vs_reference_snr_db": _snr_db(wav_fast, owav_fast) if wav_fast.size == owav_fast.size else None}

    with open(os.path.join(HERE, "summary.json"), "w") as fh:
        json.dump(summary, fh, indent=1, sort_keys=True)
    print(json.dumps(summary, indent=1, sort_keys=True))


if __name__ == "__main__":
    main()
