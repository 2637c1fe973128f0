"""ctypes binding of libunast_spectral.so (the C ABI declared in include/unast_spectral.h).

The library is the product; this module only loads it and declares argument types.  There is no
CPU fallback: if the shared object is missing this raises, and ``unast_create`` fails without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("UNAST_SPECTRAL_LIB") or os.path.join(_HERE, "libunast_spectral.so")   # override: A/B builds

OK, E_CUDA, E_ARG, E_WORKSPACE, E_DATA = 0, -1, -2, -3, -4
F32, I16 = 0, 1
PREC_F32, PREC_F64 = 0, 1
MEL_MAJOR_PER_UTT, MEL_PADDED_BTM = 0, 1

# every symbol include/unast_spectral.h declares: (name, restype, argtypes)
_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
SYMBOLS = [
    ("unast_abi_version", _i, []),
    ("unast_last_error", C.c_char_p, []),
    ("unast_create", _i, [_i, C.POINTER(_vp)]),
    ("unast_destroy", _i, [_vp]),
    ("unast_num_sms", _i, [_vp]),
    ("unast_check", _i, [_vp, _vp]),
    ("unast_get_mel_basis", _i, [_vp, _vp]),
    ("unast_trim_workspace_bytes", _sz, [_i, _i]),
    ("unast_trim_bounds", _i, [_vp, _vp, _i, _vp, _i, _i, _vp, _vp, _sz, _vp]),
    ("unast_frame_offsets", _i, [_vp, _vp, _i, _vp, _vp]),
    ("unast_frontend_workspace_bytes", _sz, [_i]),
    ("unast_frontend", _i, [_vp, _vp, _i, _vp, _vp, _vp, _i, _i, _vp, _i, _i, _vp, _vp, _sz, _vp]),
    ("unast_griffinlim_workspace_bytes", _sz, [_vp, _i, _f]),
    ("unast_griffinlim", _i, [_vp, _vp, _vp, _i, _i, _f, _vp, _f, _i, _vp, _vp, _vp, _sz, _vp]),
    ("unast_griffinlim_layout", _i, [_vp, _i, _f, _vp]),
    ("unast_griffinlim_fragment_samples", _i, [_i, _i, _i, _vp, _vp]),
    ("unast_griffinlim_fragments", _i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _i, _i, _f, _vp, _f, _vp, _vp, _sz, _vp]),
    ("unast_deemph_trim_workspace_bytes", _sz, [_i, _i]),
    ("unast_deemph_trim", _i, [_vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    ("unast_peer_alloc", _i, [_vp, _sz, C.POINTER(_vp), _vp]),
    ("unast_peer_open", _i, [_vp, _vp, C.POINTER(_vp)]),
    ("unast_peer_close", _i, [_vp, _vp]),
    ("unast_peer_free", _i, [_vp, _vp]),
    ("unast_mel_augment_workspace_bytes", _sz, [_i]),
    ("unast_mel_augment", _i, [_vp, _vp, _i, _i, _i, _vp, _vp, _i, _vp, _vp, _sz, _vp]),
    ("unast_griffinlim_set_block", _i, [_vp, _i]),
    ("unast_mel_augment_static", _i, [_vp, _vp, _i, _i, _i, _vp, _vp, _i, _vp, _vp]),
    ("unast_prenet_packed_bytes", _sz, [_i]),
    ("unast_prenet_pack", _i, [_vp, _vp, _vp, _i, _vp, _vp]),
    ("unast_mel_prenet_fc1", _i, [_vp, _vp, _i, _i, _vp, _vp, _i, _vp, _i, _i, _vp, _vp, _sz, _vp]),
    ("unast_debug_rfft", _i, [_vp, _vp, _i, _i, _vp, _vp]),
    ("unast_debug_irfft", _i, [_vp, _vp, _i, _vp, _vp]),
    ("unast_debug_pipe_peak", _i, [_vp, _i, _i, _vp, C.POINTER(C.c_double), _vp]),
]



class GlLink(C.Structure):
    """unast_gl_link of include/unast_spectral.h."""
    _fields_ = [("fragment", C.c_int32), ("side", C.c_int32), ("peer_ola", C.c_void_p * 3), ("peer_mail", C.c_void_p),
                ("my_mail", C.c_int32)]


_lib = None


class UnastError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libunast_spectral error %d: %s" % (code, msg))
        self.code = code


def load():
    """dlopen the in-tree library (built by ``unast_b200.build`` / ``__graft_entry__.build``)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "%s is missing: build it with `python -m unast_b200.build` (nvcc, sm_100a). "
            "The spectral path has no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, restype, argtypes in SYMBOLS:
        fn = getattr(lib, name)          # AttributeError here means header and library disagree
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.unast_abi_version() != 2:
        raise ImportError("libunast_spectral ABI version mismatch")
    _lib = lib
    return lib


def check(code):
    if code != OK:
        raise UnastError(code, load().unast_last_error().decode("utf-8", "replace"))
