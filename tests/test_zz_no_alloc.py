"""unast_mel_augment_static: the allocation-free variant of unast_mel_augment (scratch in a file-scope __device__ array of
the library).  It has no counterpart in the reference; it was added for callers that
cannot call cudaMalloc.  Collected last on purpose.

The test goes through the C ABI directly (ctypes), compares with the existing entry point bit for bit -- both run the same
two kernels, only the scratch differs -- on two calls in a row with different inputs (the array keeps the first call's
contents), checks that no CUDA memory was allocated by the calls, and that a batch beyond the maximum is refused."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_B = 1024                                    # UNAST_AUG_STATIC_MAX_B in include/unast_spectral.h


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def test_mel_augment_static_matches_the_workspace_entry_point(engine):
    import torch
    from unast_b200 import _lib as L
    lib, h = engine.lib, engine.h
    dev = engine.device
    g = torch.Generator(device="cpu").manual_seed(77)
    cases = []
    for (B, T, D) in ((7, 333, 80), (3, 129, 80)):                       # second call: smaller batch, stale sums beyond it
        x = torch.rand((B, T, D), generator=g).to(dev)
        f0 = torch.randint(0, T // 3, (B,), generator=g)
        t0 = torch.randint(T // 2, T - 30, (B,), generator=g)
        spans = torch.stack([f0, torch.full((B,), 17), t0, torch.full((B,), 29)], dim=1).to(torch.int32).to(dev)
        keep = (torch.rand((B, T), generator=g) > 0.3).float().to(dev)
        ref = torch.empty_like(x)
        out = torch.full_like(x, -7.0)
        nbytes = lib.unast_mel_augment_workspace_bytes(B)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        cases.append((B, T, D, x, spans, keep, ref, out, ws, nbytes))
    torch.cuda.synchronize()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    for rz in (0, 1):
        for (B, T, D, x, spans, keep, ref, out, ws, nbytes) in cases:    # two calls in a row per mode, nothing cleared between
            L.check(lib.unast_mel_augment(h, _ptr(x), B, T, D, _ptr(spans), _ptr(keep), rz, _ptr(ref), _ptr(ws), nbytes, stream))
            before = torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)
            L.check(lib.unast_mel_augment_static(h, _ptr(x), B, T, D, _ptr(spans), _ptr(keep), rz, _ptr(out), stream))
            assert (torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)) == before
            torch.cuda.synchronize()
            engine.check()
            assert torch.equal(out, ref)                                 # same kernels, same summation order: identical bits
            assert float((out - ref).abs().max()) <= 1e-6                # the tolerance of the masking tests, stated
            out.fill_(-7.0)
    # the Python wrapper takes the same path
    B, T, D, x, spans, keep, ref, out, ws, nbytes = cases[0]
    L.check(lib.unast_mel_augment(h, _ptr(x), B, T, D, _ptr(spans), _ptr(keep), 0, _ptr(ref), _ptr(ws), nbytes, stream))
    engine.mel_augment_static(x, out, spans=spans, keep=keep)
    torch.cuda.synchronize()
    assert torch.equal(out, ref)
    # no spans: no scratch needed at all, still the same result
    L.check(lib.unast_mel_augment(h, _ptr(x), B, T, D, None, _ptr(keep), 0, _ptr(ref), _ptr(ws), nbytes, stream))
    L.check(lib.unast_mel_augment_static(h, _ptr(x), B, T, D, None, _ptr(keep), 0, _ptr(out), stream))
    torch.cuda.synchronize()
    assert torch.equal(out, ref)


def test_mel_augment_static_refuses_what_does_not_fit(engine):
    import torch
    from unast_b200 import _lib as L
    lib, h = engine.lib, engine.h
    x = torch.zeros((MAX_B + 1, 4, 80), device=engine.device)
    out = torch.empty_like(x)
    spans = torch.zeros((MAX_B + 1, 4), dtype=torch.int32, device=engine.device)
    rc = lib.unast_mel_augment_static(h, _ptr(x), MAX_B + 1, 4, 80, _ptr(spans), None, 0, _ptr(out), ctypes.c_void_p(0))
    assert rc != 0
    with pytest.raises(Exception, match="UNAST_AUG_STATIC_MAX_B"):
        L.check(rc)
    # the largest batch that fits runs
    L.check(lib.unast_mel_augment_static(h, _ptr(x), MAX_B, 4, 80, _ptr(spans), None, 0, _ptr(out), ctypes.c_void_p(0)))
    torch.cuda.synchronize()
    engine.check()
    assert float(out[:MAX_B].abs().max()) == 0.0
