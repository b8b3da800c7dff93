"""ctypes binding of libfm_b200.so (include/fm_b200.h).

The library is the product: there is no Python/PyTorch implementation of the hot path behind
it and no CPU fallback.  If the shared object is missing or a call fails, this module raises.
"""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# FM_B200_LIB: developer switch for A/B builds of the same library (tools/ab_build.sh); the product is libfm_b200.so
LIB_PATH = os.environ.get("FM_B200_LIB") or os.path.join(HERE, "libfm_b200.so")

FM_OK = 0
DTYPE_BF16, DTYPE_F32 = 0, 1
STATUS_WORDS = 8
RECORD_BYTES = 40
MAX_CHUNKS = 64
ST_COUNT, ST_FLAGS, ST_MAX_CAND, ST_LOST, ST_CAND_TOTAL, ST_REDO = 0, 1, 2, 3, 4, 5
FLAG_CAND_OVERFLOW, FLAG_OUT_OVERFLOW, FLAG_TIE_OVERFLOW, FLAG_RANGE = 1, 2, 4, 8
POSE_POLISH, POSE_IDENTITY_K = 1, 2
POSE_STATIC_SCRATCH_BYTES = 2 * 1024 * 1024

_c = ctypes
_P, _I, _F, _Z = _c.c_void_p, _c.c_int, _c.c_float, _c.c_size_t

# name -> (restype, argtypes); mirrors include/fm_b200.h declaration by declaration
SIGNATURES = {
    "fm_version": (_I, []),
    "fm_last_error_string": (_c.c_char_p, []),
    "fm_device_sm_count": (_I, []),
    "fm_launch_count": (_c.c_longlong, []),
    "fm_coarse_workspace_bytes": (_Z, [_I] * 6),
    "fm_coarse_match": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _F, _F, _I, _F, _P, _P,
                             _P, _P, _P, _P, _P, _P, _P, _I, _P, _Z, _I, _P]),
    "fm_debug_coarse_match_simt": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _F, _F, _I, _F,
                                        _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _P, _Z, _I, _P]),
    "fm_debug_coarse_lse": (_I, [_P, _I, _I, _I, _I, _I, _I, _P, _P, _P]),
    "fm_profile_coarse_stats": (_I, [_P, _P, _I, _I, _I, _I, _F, _F, _P, _Z, _I, _I,
                                     _c.POINTER(_c.c_float), _P]),
    "fm_fine_to_nhwc_bf16": (_I, [_P, _I, _I, _I, _I, _I, _P, _P]),
    "fm_fine_gather_workspace_bytes": (_Z, [_I] * 6),
    "fm_fine_gather": (_I, [_P, _P, _P, _P, _I, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I,
                            _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _I, _P, _Z, _P]),
    "fm_pack_matches": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _P, _P]),
    "fm_pack_matches_append": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _P, _I, _I, _P, _P]),
    "fm_profile_enable": (_I, [_I]),
    "fm_profile_collect": (_I, [_c.POINTER(_c.c_float), _c.POINTER(_c.c_int)]),
    "fm_fine_transformer_workspace_bytes": (_Z, [_I]),
    "fm_fine_transformer": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _Z, _P]),
    "fm_coarse_transformer_workspace_bytes": (_Z, [_I, _I, _I, _I]),
    "fm_coarse_transformer": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _P, _Z, _P]),
    "fm_backbone_emit": (_I, [_P, _c.c_longlong, _I, _I, _P, _P, _F, _P, _P, _P, _P]),
    "fm_pose_workspace_bytes": (_Z, [_I, _I, _I]),
    "fm_pose_from_matches": (_I, [_P, _P, _P, _I, _P, _I, _I, _P, _F, _I, _I, _c.c_uint32, _P, _P, _P, _P, _P, _P, _P,
                                  _Z, _P]),
    "fm_debug_pose_valid_hypotheses": (_I, [_P, _I, _I, _I, _P, _P]),
    "fm_pose_from_matches_static": (_I, [_P, _P, _P, _I, _P, _I, _I, _P, _F, _I, _I, _c.c_uint32, _P, _P, _P, _P, _P, _P,
                                         _P]),
    "fm_fine_match": (_I, [_P, _P, _I, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _F, _P, _P, _P]),
}

_lib = None


class FmError(RuntimeError):
    pass


def load():
    """Loads libfm_b200.so (built by featurematching_b200.build / __graft_entry__.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FmError(f"{LIB_PATH} not found: run `python -m featurematching_b200.build` "
                      "(the matching head has no fallback implementation)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what):
    if rc != FM_OK:
        msg = load().fm_last_error_string().decode("utf-8", "replace")
        raise FmError(f"{what} failed with code {rc}: {msg}")


def ptr(t):
    """Device pointer of a torch tensor (or None) as a void*."""
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def dtype_code(t):
    import torch
    if t.dtype == torch.bfloat16:
        return DTYPE_BF16
    if t.dtype == torch.float32:
        return DTYPE_F32
    raise FmError(f"unsupported dtype {t.dtype}: features must be bfloat16 or float32")


def stream_ptr(device):
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
