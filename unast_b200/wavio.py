"""Minimal WAV / NPY I/O for the spectral path (no soundfile / librosa in this image).

``read_wav`` plays the role of ``librosa.load(fpath, sr=22050)`` at reference src/utils.py:245 for
files that already are 22,050 Hz (LJSpeech, samples/test.wav): PCM16 data is returned as int16 so
that the /32768 conversion happens on the GPU inside the front-end's load (fused), PCM32 / float32
are returned as float32.  Resampling is out of scope and raises instead of silently differing.
"""
from __future__ import annotations

import struct

import numpy as np

from . import audio_parameters as ap


def read_wav(path, expect_sr=ap.sr):
    with open(path, "rb") as fh:
        b = fh.read()
    if len(b) < 12 or b[:4] != b"RIFF" or b[8:12] != b"WAVE":
        raise ValueError("%s: not a RIFF/WAVE file" % path)
    i, fmt, data = 12, None, None
    while i + 8 <= len(b):
        cid, sz = b[i:i + 4], struct.unpack("<I", b[i + 4:i + 8])[0]
        if cid == b"fmt ":
            fmt = struct.unpack("<HHIIHH", b[i + 8:i + 24])
        elif cid == b"data":
            data = b[i + 8:i + 8 + sz]
        i += 8 + sz + (sz & 1)
    if fmt is None or data is None:
        raise ValueError("%s: missing fmt or data chunk" % path)
    tag, nch, rate, _, _, bits = fmt
    if tag == 0xFFFE and len(b) >= 44:          # WAVE_FORMAT_EXTENSIBLE: sub-format code follows
        tag = 3 if bits == 32 and b"\x03\x00\x00\x00\x00\x00\x10\x00" in b[:80] else 1
    if expect_sr is not None and rate != expect_sr:
        raise ValueError("%s: sample rate %d != %d; resampling (librosa kaiser_best) is out of scope" % (path, rate, expect_sr))
    if tag == 1 and bits == 16:
        y = np.frombuffer(data, "<i2")
    elif tag == 1 and bits == 32:
        y = (np.frombuffer(data, "<i4").astype(np.float64) / 2147483648.0).astype(np.float32)
    elif tag == 3 and bits == 32:
        y = np.frombuffer(data, "<f4")
    else:
        raise ValueError("%s: unsupported WAV encoding (tag %d, %d bits)" % (path, tag, bits))
    if nch > 1:                                  # librosa.to_mono
        y = y.reshape(-1, nch).astype(np.float32)
        if tag == 1 and bits == 16:
            y = y / np.float32(32768.0)
        y = y.mean(axis=1).astype(np.float32)
    return np.ascontiguousarray(y), rate


def wav_info(path, expect_sr=ap.sr):
    """(samples per channel, numpy dtype of read_wav's result, channels) from the header alone (no payload read)."""
    with open(path, "rb") as fh:
        head = fh.read(12)
        if len(head) < 12 or head[:4] != b"RIFF" or head[8:12] != b"WAVE":
            raise ValueError("%s: not a RIFF/WAVE file" % path)
        fmt = None
        while True:
            ch = fh.read(8)
            if len(ch) < 8:
                raise ValueError("%s: missing fmt or data chunk" % path)
            cid, sz = ch[:4], struct.unpack("<I", ch[4:])[0]
            if cid == b"fmt ":
                body = fh.read(sz + (sz & 1))
                fmt = struct.unpack("<HHIIHH", body[:16])
            elif cid == b"data":
                if fmt is None:
                    raise ValueError("%s: data chunk before fmt" % path)
                tag, nch, rate, _, align, bits = fmt
                if expect_sr is not None and rate != expect_sr:
                    raise ValueError("%s: sample rate %d != %d; resampling (librosa kaiser_best) is out of scope" % (path, rate, expect_sr))
                dt = np.int16 if (tag == 1 and bits == 16 and nch == 1) else np.float32
                return sz // max(1, align), dt, nch
            else:
                fh.seek(sz + (sz & 1), 1)


def npy_header(shape, fortran_order):
    """The bytes np.save writes before a float32 payload of this shape / order (format 1.0, numpy's own writer)."""
    import io
    buf = io.BytesIO()
    np.lib.format.write_array_header_1_0(buf, {"descr": "<f4", "fortran_order": bool(fortran_order), "shape": tuple(int(v) for v in shape)})
    return buf.getvalue()


def write_wav_pcm16(path, y, rate=ap.sr):
    """16-bit PCM WAV, what ``sf.write(path, wav, sr)`` (gl_vocoder.py:34) produces: python-soundfile's default subtype
    for ``.wav`` is PCM_16 whatever the array dtype, and libsndfile converts floats (normalised, clipping off -- the
    defaults) as ``lrint(x * 0x7FFF)`` truncated to 16 bits (libsndfile src/pcm.c, d2les_array: round half to even, values
    outside [-1, 1] wrap).  libsndfile is not installable in this image, so this restates its published source; no test
    here pins it against the real library."""
    v = np.rint(np.asarray(y, dtype=np.float64) * 32767.0).astype(np.int64)
    data = (v & 0xFFFF).astype("<u2").tobytes()
    with open(path, "wb") as fh:
        fh.write(b"RIFF" + struct.pack("<I", 36 + len(data)) + b"WAVE")
        fh.write(b"fmt " + struct.pack("<IHHIIHH", 16, 1, 1, rate, rate * 2, 2, 16))
        fh.write(b"data" + struct.pack("<I", len(data)) + data)


def write_wav_float32(path, y, rate=ap.sr):
    """IEEE float32 WAV (``sf.write(..., subtype='FLOAT')`` / ``scipy.io.wavfile.write`` as Transformer-TTS
    synthesis.py:50 uses): opt-in, not what gl_vocoder.py:34 writes."""
    y = np.ascontiguousarray(np.asarray(y, dtype="<f4"))
    data = y.tobytes()
    with open(path, "wb") as fh:
        fh.write(b"RIFF" + struct.pack("<I", 4 + 26 + 12 + 8 + len(data)) + b"WAVE")
        fh.write(b"fmt " + struct.pack("<IHHIIHHH", 18, 3, 1, rate, rate * 4, 4, 32, 0))
        fh.write(b"fact" + struct.pack("<II", 4, y.size))
        fh.write(b"data" + struct.pack("<I", len(data)) + data)
