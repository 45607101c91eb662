"""ctypes loader for libcabac_b200.so (built in-tree by __graft_entry__.build()).

Fails loudly when the CUDA library is missing: there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CABAC_B200_LIB") or os.path.join(HERE, "libcabac_b200.so")  # env: dev A/B builds

_lib = None

_P = C.c_void_p
_SZ = C.c_size_t

_SIGS = {
    "cabac_create": (_P, [C.c_int]),
    "cabac_destroy": (None, [_P]),
    "cabac_set_stream": (C.c_int, [_P, _P]),
    "cabac_sync": (C.c_int, [_P]),
    "cabac_set_encoder_mode": (C.c_int, [_P, C.c_int]),
    "cabac_set_decoder_mode": (C.c_int, [_P, C.c_int]),
    "cabac_last_error": (C.c_char_p, [_P]),
    "cabac_version": (C.c_char_p, []),
    "cabac_launch_count": (C.c_uint64, [_P]),
    "cabac_host_alloc": (_P, [_SZ]),
    "cabac_host_free": (None, [_P]),
    "cabac_encode_residual_packed": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _SZ, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P]),
    "cabac_ctx_init": (C.c_int, [_P, _P, _SZ, _P]),
    "cabac_ctx_init_dev": (C.c_int, [_P, _P, _SZ, _P]),
    "cabac_ctx_rates": (C.c_int, [_P]),
    "cabac_ctx_init_table": (C.c_int, [C.c_int, _P]),
    "cabac_encode_bins": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_encode_bins_dev": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_encode_stream": (C.c_int, [_P, _P, _P, _P, _P, _P, _SZ, _P, _SZ, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_encode_stream_dev": (C.c_int, [_P, _P, _P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_decode_bins": (C.c_int, [_P, _P, _P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P]),
    "cabac_decode_bins_dev": (C.c_int, [_P, _P, _P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P]),
    "cabac_encode_residual": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _SZ, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_encode_residual_dev": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_decode_residual": (C.c_int, [_P, _P, _P, _P, _P, _P, _SZ, _P, _P, _SZ, _P, _P, _P, _P, _P]),
    "cabac_decode_residual_dev": (C.c_int, [_P, _P, _P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_decode_residual_static_dev": (C.c_int, [_P, _P, _P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_estimate_residual_bits": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _SZ, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_estimate_residual_bits_dev": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_residual_ctx_bins": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _SZ, _P, _P, _P, _P, _P]),
    "cabac_residual_ctx_bins_dev": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P]),
    "cabac_ctx_init_id": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _P]),
    "cabac_ctx_init_id_dev": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _P]),
    "cabac_frac_bits_table": (C.c_int, [_P]),
    "cabac_count_emulations": (C.c_int, [_P, _P, _P, _SZ, _P]),
    "cabac_count_emulations_dev": (C.c_int, [_P, _P, _P, _SZ, _P]),
    "cabac_insert_emulation_prevention": (C.c_int, [_P, _P, _SZ, _P, _SZ, _P]),
    "cabac_insert_emulation_prevention_dev": (C.c_int, [_P, _P, _SZ, _P, _SZ, _P]),
    "cabac_encode_residual_wpp": (C.c_int, [_P, _P, _P, _P, _P, _SZ, _P, _SZ, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P]),
    "cabac_encode_residual_wpp_dev": (C.c_int, [_P, _P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_decode_residual_wpp": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _SZ, _P, _P, _SZ, _P, _P, _P, _P, _P]),
    "cabac_decode_residual_wpp_dev": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_encode_stream_wpp": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _SZ, _P, _SZ, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_encode_stream_wpp_dev": (C.c_int, [_P, _P, _P, _P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_remove_emulation_prevention": (C.c_int, [_P, _P, _SZ, C.c_int, _P, _P, _P, _P]),
    "cabac_remove_emulation_prevention_dev": (C.c_int, [_P, _P, _SZ, C.c_int, _P, _P, _P, _P]),
    "cabac_entry_point_syntax": (C.c_int, [_P, _P, _P, _SZ, _P, _P]),
    "cabac_entry_point_syntax_dev": (C.c_int, [_P, _P, _P, _SZ, _P, _P]),
    "cabac_substream_ranges": (C.c_int, [_P, _P, _SZ, _P, _SZ, C.c_uint64, C.c_uint64, _P]),
    "cabac_substream_ranges_dev": (C.c_int, [_P, _P, _SZ, _P, _SZ, C.c_uint64, C.c_uint64, _P]),
    "cabac_pin_descriptors": (C.c_int, [_P, _P, _P, _P, _SZ]),
    "cabac_unpin_descriptors": (C.c_int, [_P]),
    "cabac_encode_residual_i16_dev": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_decode_residual_i16_dev": (C.c_int, [_P, _P, _P, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P, _P]),
    "cabac_encode_residual_packed_i16": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _SZ, _P, _P, _P, _SZ, _P, _P, _P, _P, _P, _P]),
    "cabac_decode_residual_i16": (C.c_int, [_P, _P, _P, _P, _P, _P, _SZ, _P, _P, _SZ, _P, _P, _P, _P, _P]),
    "cabac_decode_slices": (C.c_int, [_P, _P, _P, _P, _P, _SZ] + [_P] * 15),
    "cabac_decode_slices_dev": (C.c_int, [_P, _P, _P, _P, _P, _SZ] + [_P] * 15),
    "cabac_encode_slices": (C.c_int, [_P, _P, _P, _SZ] + [_P] * 8 + [_SZ, _P, _SZ] + [_P] * 6),
    "cabac_encode_slices_dev": (C.c_int, [_P, _P, _P, _SZ] + [_P] * 10 + [_SZ] + [_P] * 6),
    "cabac_encode_slices_wpp": (C.c_int, [_P, _P, _P, _SZ] + [_P] * 8 + [_SZ, _P, _SZ, _P, _SZ] + [_P] * 6),
    "cabac_encode_slices_wpp_dev": (C.c_int, [_P, _P, _P, _SZ] + [_P] * 10 + [_SZ, _P, _SZ] + [_P] * 6),
    "cabac_pack": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _SZ, _P]),
    "cabac_pack_dev": (C.c_int, [_P, _P, _P, _P, _SZ, _P, _P]),
}

EXPORTS = tuple(_SIGS)


def load() -> C.CDLL:
    """Returns the loaded library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
                "entropy_coding_b200 has no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
