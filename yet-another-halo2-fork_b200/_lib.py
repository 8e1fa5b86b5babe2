"""ctypes binding of libh2b200.so (include/h2b200.h).  Fails loudly if the CUDA library is missing:
there is no CPU fallback anywhere in this package."""
from __future__ import annotations

import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libh2b200.so")

H2B_OK, H2B_E_INVALID, H2B_E_CUDA, H2B_E_NOMEM, H2B_E_NODEVICE, H2B_E_CONSTRAINT = 0, -1, -2, -3, -4, -5
MSM_WINDOW_BYTES = 128
MSM_MAX_WINDOWS = 64
PEER_HANDLE_BYTES = 64

# every symbol include/h2b200.h declares: name -> (restype, argtypes)
_vp, _sz, _u32, _u64, _int = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint32, ctypes.c_uint64, ctypes.c_int
_pu32, _pu64, _pd = ctypes.POINTER(ctypes.c_uint32), ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_double)
SYMBOLS = {
    "h2b_init": (_int, [ctypes.POINTER(_vp), ctypes.POINTER(_int), _int]),
    "h2b_destroy": (None, [_vp]),
    "h2b_last_error": (ctypes.c_char_p, [_vp]),
    "h2b_device_count": (_int, []),
    "h2b_version": (ctypes.c_char_p, []),
    "h2b_last_notice": (ctypes.c_char_p, [_vp]),
    "h2b_set_option": (_int, [_vp, ctypes.c_char_p, ctypes.c_char_p]),
    "h2b_bases_info": (_int, [_vp, _u64, _pu32, _pu64]),
    "h2b_peer_export": (_int, [_vp, _vp, ctypes.POINTER(_vp)]),
    "h2b_peer_attach": (_int, [_vp, _int, _int, _vp, ctypes.POINTER(_vp)]),
    "h2b_msm_sharded": (_int, [_vp, _u64, _vp, _int, _sz, _sz, _vp]),
    "h2b_bases_register": (_int, [_vp, _vp, _sz, _pu64]),
    "h2b_bases_register_device": (_int, [_vp, _vp, _sz, _pu64]),
    "h2b_bases_release": (_int, [_vp, _u64]),
    "h2b_msm": (_int, [_vp, _u64, _vp, _sz, _vp]),
    "h2b_msm_raw": (_int, [_vp, _vp, _vp, _sz, _vp]),
    "h2b_msm_device": (_int, [_vp, _u64, _vp, _sz, _vp]),
    "h2b_msm_windows": (_int, [_vp, _u64, _vp, _int, _sz, _sz, _vp, _pu32, _pu32]),
    "h2b_msm_batch": (_int, [_vp, _u64, ctypes.POINTER(_vp), _sz, _sz, _vp]),
    "h2b_msm_batch_device": (_int, [_vp, _u64, ctypes.POINTER(_vp), _sz, _sz, _vp]),
    "h2b_msm_combine": (_int, [_vp, _u32, _u32, _u32, _vp]),
    "h2b_ntt": (_int, [_vp, _vp, _u32, _vp, _vp]),
    "h2b_ntt_device": (_int, [_vp, _vp, _u32, _vp, _vp]),
    "h2b_coeff_to_extended": (_int, [_vp, _vp, _vp, _u32, _u32, _vp, _vp, _vp]),
    "h2b_coeff_to_extended_device": (_int, [_vp, _vp, _vp, _u32, _u32, _vp, _vp, _vp]),
    "h2b_extended_to_coeff": (_int, [_vp, _vp, _vp, _u32, _vp, _vp, _vp, _vp, _sz, _vp, _sz]),
    "h2b_extended_to_coeff_device": (_int, [_vp, _vp, _vp, _u32, _vp, _vp, _vp, _vp, _sz, _vp, _sz]),
    "h2b_ntt_batch": (_int, [_vp, ctypes.POINTER(_vp), _sz, _u32, _vp, _vp]),
    "h2b_coeff_to_extended_batch": (_int, [_vp, ctypes.POINTER(_vp), ctypes.POINTER(_vp), _sz, _u32, _u32, _vp, _vp, _vp]),
    "h2b_coeff_to_extended_batch_device": (_int, [_vp, ctypes.POINTER(_vp), ctypes.POINTER(_vp), _sz, _u32, _u32, _vp, _vp, _vp]),
    "h2b_batch_invert": (_int, [_vp, _vp, _sz]),
    "h2b_batch_invert_device": (_int, [_vp, _vp, _sz]),
    "h2b_batch_invert_device_noalloc": (_int, [_vp, _sz, _vp, ctypes.POINTER(ctypes.c_char_p)]),
    "h2b_fr_mul_device": (_int, [_vp, _vp, _vp, _vp, _vp, _sz]),
    "h2b_grand_product": (_int, [_vp, _vp, _sz, _vp, _vp, _sz]),
    "h2b_grand_product_device": (_int, [_vp, _vp, _sz, _vp, _vp, _sz]),
    "h2b_eval_polynomial": (_int, [_vp, _vp, _sz, _vp, _vp]),
    "h2b_eval_polynomial_device": (_int, [_vp, _vp, _sz, _vp, _vp]),
    "h2b_eval_polynomial_batch_device": (_int, [_vp, _vp, _sz, _vp, _sz, _vp]),
    "h2b_kate_division": (_int, [_vp, _vp, _sz, _vp, _vp]),
    "h2b_kate_division_device": (_int, [_vp, _vp, _sz, _vp, _vp]),
    "h2b_quotient_graph_device": (_int, [_vp, _vp, ctypes.POINTER(_vp), _sz, ctypes.POINTER(_vp), _sz, ctypes.POINTER(_vp), _sz,
                                         _vp, _sz, _vp, _vp, _vp, _vp, _vp, _vp, _u32, _u32]),
    "h2b_quotient_permutation_device": (_int, [_vp, _vp, ctypes.POINTER(_vp), _sz, ctypes.POINTER(_vp), ctypes.POINTER(_vp), _sz, _sz,
                                               _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, ctypes.c_int32, _u32, _u32]),
    "h2b_quotient_lookup_device": (_int, [_vp] * 12 + [_u32, _u32]),
    "h2b_quotient_shuffle_device": (_int, [_vp] * 9 + [_u32, _u32]),
    "h2b_last_call_ms": (ctypes.c_double, [_vp]),
    "h2b_profile_enable": (_int, [_vp, _int]),
    "h2b_profile_read": (_int, [_vp, ctypes.c_char_p, _sz]),
    "h2b_launch_count": (_u64, [_vp]),
    "h2b_imad_peak": (_int, [_vp, _pd, _pd]),
    "h2b_synth_bases": (_int, [_vp, _vp, _sz, _u64]),
    "h2b_g1_fixed_base_mul_device": (_int, [_vp, _vp, _sz, _vp]),
    "h2b_fr_from_u512_device": (_int, [_vp, _vp, _sz, _vp]),
    "h2b_lookup_permute_device": (_int, [_vp, _vp, _vp, _sz, _sz, _vp, _vp]),
    "h2b_fr_random_chacha20_device": (_int, [_vp, _vp, ctypes.POINTER(ctypes.c_uint64), _sz, _vp]),
}



class Graph(ctypes.Structure):
    """`h2b_graph` of include/h2b200.h."""
    _fields_ = [("program", ctypes.POINTER(ctypes.c_uint32)), ("n_words", _sz), ("constants32", _vp), ("n_constants", _sz),
                ("rotations", ctypes.POINTER(ctypes.c_int32)), ("n_rotations", _sz), ("n_intermediates", _u32)]


_lib = None


class H2bError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"h2b200 error {code}: {msg}")
        self.code = code


def load() -> ctypes.CDLL:
    """Load libh2b200.so and bind every declared symbol.  Raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m __graft_entry__` or "
            "`python yet-another-halo2-fork_b200/build.py` (nvcc, sm_100a).  There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)       # AttributeError here = header / library drift
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
