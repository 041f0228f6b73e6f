"""ctypes binding of include/alexcap.h (libalexcap_b200.so).  No torch types cross this boundary: only raw
device pointers, sizes and a cudaStream_t.  There is NO fallback: if the library is missing or the device is not
sm_100 every entry point raises."""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

ALEXCAP_MAX_LAYERS = 8
F32, BF16 = 0, 1
OPT_MAX_TENSORS = 48      # ALEXCAP_OPT_MAX_TENSORS (include/alexcap.h)


class Dims(C.Structure):
    _fields_ = [('B', C.c_int32), ('L', C.c_int32), ('D', C.c_int32), ('H', C.c_int32), ('E', C.c_int32),
                ('Vp', C.c_int32), ('S', C.c_int32), ('num_layers', C.c_int32), ('dtype', C.c_int32),
                ('need_dfeat', C.c_int32), ('ignore_index', C.c_int32), ('label_smoothing', C.c_float),
                ('reg_weight', C.c_float), ('dropout_p', C.c_float), ('seed', C.c_uint64), ('offset', C.c_uint64),
                ('philox_state', C.c_void_p)]


ATTN_FIELDS = ['init_h_w', 'init_h_b', 'init_c_w', 'init_c_b', 'f_beta_w', 'f_beta_b', 'deep_w', 'deep_b',
               'att_U_w', 'att_U_b', 'att_W_w', 'att_W_b', 'att_v_w', 'att_v_b', 'emb_w',
               'lstm_w_ih', 'lstm_w_hh', 'lstm_b_ih', 'lstm_b_hh']
# C field -> reference state_dict name (AlexCap/AttentionLanguageModule.py:46-58)
ATTN_STATE_NAMES = ['init_h.weight', 'init_h.bias', 'init_c.weight', 'init_c.bias', 'f_beta.weight', 'f_beta.bias',
                    'deep_output.weight', 'deep_output.bias', 'attention.U.weight', 'attention.U.bias',
                    'attention.W.weight', 'attention.W.bias', 'attention.v.weight', 'attention.v.bias',
                    'embedding.weight', 'lstm.weight_ih', 'lstm.weight_hh', 'lstm.bias_ih', 'lstm.bias_hh']


class AttnParams(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ATTN_FIELDS]


class LstmParams(C.Structure):
    _fields_ = [('enc_w', C.c_void_p), ('enc_b', C.c_void_p), ('emb_w', C.c_void_p), ('out_w', C.c_void_p),
                ('out_b', C.c_void_p),
                ('w_ih', C.c_void_p * ALEXCAP_MAX_LAYERS), ('w_hh', C.c_void_p * ALEXCAP_MAX_LAYERS),
                ('b_ih', C.c_void_p * ALEXCAP_MAX_LAYERS), ('b_hh', C.c_void_p * ALEXCAP_MAX_LAYERS)]


_P, _I32, _I64, _F, _SZ = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_size_t
_DP, _AP, _LP = C.POINTER(Dims), C.POINTER(AttnParams), C.POINTER(LstmParams)

# name -> (restype, argtypes): one entry per prototype in include/alexcap.h
PROTOTYPES = {
    'alexcap_abi_version': (C.c_int, []),
    'alexcap_last_error': (C.c_char_p, []),
    'alexcap_device_check': (C.c_int, []),
    'alexcap_launch_count': (C.c_ulonglong, []),
    'alexcap_set_gemm_engine': (C.c_int, [C.c_int]),
    'alexcap_get_gemm_engine': (C.c_int, []),
    'alexcap_attn_workspace_bytes': (_SZ, [_DP]),
    'alexcap_attn_workspace_region': (C.c_int, [_DP, C.c_char_p, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]),
    'alexcap_set_time_loop': (C.c_int, [_I32]),
    'alexcap_time_loop_profile': (C.c_int, [C.POINTER(C.c_ulonglong), _I32]),
    'alexcap_mean_over_L': (C.c_int, [_P, _I32, _I32, _I32, _P, _P]),
    'alexcap_count_valid_targets': (C.c_int, [_P, _I32, _I32, _P, _P]),
    'alexcap_attn_train_fused': (C.c_int, [_DP, _AP, _P, _P, _AP, _P, _P, _P, _P, _P, _P, _SZ, _P]),
    'alexcap_clip_adam_scratch_floats': (C.c_size_t, []),
    'alexcap_clip_adam_step': (C.c_int, [C.c_int32, _P, _P, _P, _P, _P, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float,
                                         C.c_float, C.c_int64, _P, _P, _P]),
    'alexcap_features_from_nchw': (C.c_int, [C.c_int32, C.c_int32, C.c_int32, _P, _P, _P]),
    'alexcap_dfeat_to_nchw': (C.c_int, [C.c_int32, C.c_int32, C.c_int32, _P, _P, C.c_int32, _P]),
    'alexcap_attn_train_phase': (C.c_int, [_DP, _AP, _P, _P, _AP, _P, _P, _P, _P, _P, _P, _SZ, C.c_int32, _P]),
    'alexcap_attn_forward': (C.c_int, [_DP, _AP, _P, _P, _P, _P, _P, _P, _SZ, _P]),
    'alexcap_attn_backward': (C.c_int, [_DP, _AP, _P, _P, _P, _AP, _P, _P, _SZ, _P]),
    'alexcap_lstm_workspace_bytes': (_SZ, [_DP]),
    'alexcap_lstm_train_fused': (C.c_int, [_DP, _LP, _P, _P, _LP, _P, _P, _P, _P, _P, _SZ, _P]),
    'alexcap_lstm_forward': (C.c_int, [_DP, _LP, _P, _P, _P, _P, _P, _SZ, _P]),
    'alexcap_lstm_backward': (C.c_int, [_DP, _LP, _P, _P, _LP, _P, _P, _SZ, _P]),
    'alexcap_make_target': (C.c_int, [_P, _P, _P, _I32, _I32, _I32, _I32, _P]),
    'alexcap_cross_entropy': (C.c_int, [_P, _P, _I32, _I32, _I64, _F, _I32, _P, _P, _P, _P]),
    'alexcap_cross_entropy_static_max_rows': (C.c_int32, []),
    'alexcap_cross_entropy_static': (C.c_int, [_P, _P, _I32, _I32, _I64, _F, _I32, _P, _P, _P]),
    'alexcap_attn_decode_workspace_bytes': (_SZ, [_DP, _I32]),
    'alexcap_attn_decode_greedy': (C.c_int, [_DP, _AP, _P, _P, _P, _P, _SZ, _P]),
    'alexcap_attn_decode_beam': (C.c_int, [_DP, _AP, _P, _I32, _P, _P, _P, _P, _P, _SZ, _P]),
    'alexcap_lstm_decode_workspace_bytes': (_SZ, [_DP]),
    'alexcap_lstm_decode_greedy': (C.c_int, [_DP, _LP, _P, _P, _P, _SZ, _P]),
    'alexcap_lstm_decode_beam_workspace_bytes': (_SZ, [_DP, _I32]),
    'alexcap_lstm_decode_beam': (C.c_int, [_DP, _LP, _P, _I32, _P, _P, _P, _P, _SZ, _P]),
    'alexcap_gemm_f32': (C.c_int, [_I32, _I32, _I32, _I32, _I32, _F, _P, _I64, _P, _I64, _P, _I64, _P, _P, _I32, _I32,
                                   _I32, _P]),
    'alexcap_attn_step_fwd': (C.c_int, [_I32, _I32, _I32, _I32, _P, _P, _P, _I64, _P, _I64, _P, _P, _I64, _P, _P, _P,
                                        _I64, _I32, _P]),
    'alexcap_attn_step_bwd': (C.c_int, [_I32, _I32, _I32, _I32, _P, _P, _P, _I64, _P, _P, _I64, _P, _P, _P, _I64, _P,
                                        _I64, _P, _P, _I64, _P, _I64, _P, _I64, _P, _P, _I32, _P]),
}

_lib = None


class AlexcapError(RuntimeError):
    pass


def lib_path() -> str:
    return _build.LIB_PATH


def load(build_if_missing: bool = True):
    """Load (building in-tree if needed) libalexcap_b200.so and bind every prototype of include/alexcap.h."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        if not build_if_missing:
            raise AlexcapError(f'{path} missing: run `python -m imagecaptioning_b200.build`')
        _build.build_library()
    lib = C.CDLL(path)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)      # AttributeError here == header/ABI mismatch: fail loudly
        fn.restype, fn.argtypes = res, args
    if lib.alexcap_abi_version() != 2:
        raise AlexcapError('libalexcap_b200.so ABI version mismatch')
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().alexcap_last_error().decode('utf-8', 'replace')
        raise AlexcapError(f'{what} failed (code {rc}): {msg}')
