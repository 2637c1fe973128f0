"""Generates tests/golden/augment.npz by running the REFERENCE'S OWN ``specaugment`` and ``noise_fn``.

Run in the build container only (needs /root/reference):  python tests/golden/make_augment_golden.py

/root/reference/src/utils.py is imported unmodified through the same stand-in modules as make_golden.py (its
top-level imports of librosa / soundfile / jiwer are not installed here); the two functions themselves
(utils.py:40-49, :51-75) need torch and ``random`` only, so what the fixture holds IS the reference's output:

  mel [B, T, 80]   seeded input batch (zero padded past mel_len, like collate_fn_transformer's)
  mel_len [B]
  sa_mean / sa_zero   utils.specaugment(mel, mel_len) / (..., replace_with_zero=True) after
                      torch.manual_seed(SEED); random.seed(SEED)
  noise, noise_mask   utils.noise_fn(mel) after torch.manual_seed(SEED), and the mask recovered from the same seed
  short_error         the exception text the reference raises for an utterance shorter than its drawn span
"""
from __future__ import annotations

import os
import random
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
SEED = 1234


def main():
    import make_golden
    make_golden._install_stubs()
    sys.path.insert(0, make_golden.REF_SRC)
    import utils as ref_utils            # /root/reference/src/utils.py, unmodified

    rs = np.random.RandomState(7)
    B, T, D = 6, 160, 80
    mel_len = np.array([160, 151, 140, 122, 97, 64], dtype=np.int64)
    mel = rs.uniform(0.05, 0.95, size=(B, T, D)).astype(np.float32)
    for i, n in enumerate(mel_len):
        mel[i, n:] = 0.0
    tm, tl = torch.from_numpy(mel), torch.from_numpy(mel_len)

    out = {"mel": mel, "mel_len": mel_len, "seed": np.int64(SEED)}
    for key, rz in (("sa_mean", False), ("sa_zero", True)):
        torch.manual_seed(SEED)
        random.seed(SEED)
        out[key] = ref_utils.specaugment(tm, tl, replace_with_zero=rz).numpy()
        assert not np.shares_memory(out[key], mel) and np.array_equal(tm.numpy(), mel)     # input untouched
    torch.manual_seed(SEED)
    out["noise"] = ref_utils.noise_fn(tm).numpy()
    torch.manual_seed(SEED)
    gen = torch.zeros((B, T)).fill_(1 - .3)
    out["noise_mask"] = torch.bernoulli(gen).numpy()
    assert np.array_equal(out["noise"], mel * out["noise_mask"][..., None])

    # error behaviour: a 10-frame utterance cannot hold a drawn span of more frames
    torch.manual_seed(SEED)
    random.seed(SEED)
    try:
        ref_utils.specaugment(tm[:, :10].contiguous(), torch.full((B,), 10, dtype=torch.int64))
        msg = ""
    except Exception as e:                                                                # noqa: BLE001
        msg = "%s: %s" % (type(e).__name__, e)
    out["short_error"] = np.array(msg)
    np.savez_compressed(os.path.join(HERE, "augment.npz"), **out)
    print({k: (v.shape if hasattr(v, "shape") else v) for k, v in out.items()}, msg)


if __name__ == "__main__":
    main()
