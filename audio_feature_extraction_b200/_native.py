"""ctypes binding of libafx_b200.so (C ABI: include/afx_b200.h).

There is no CPU fallback: if the CUDA library has not been built the import of this module's
`lib()` raises, and every compute call needs a B200 (sm_100a) device.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_PKG = Path(__file__).resolve().parent
LIB_PATH = _PKG / "lib" / "libafx_b200.so"

AFX_FEAT_F0, AFX_FEAT_MFCC, AFX_FEAT_ENERGY, AFX_FEAT_ALL = 0x1, 0x2, 0x4, 0x7
AFX_FEAT_PREPROCESSED = 0x100   # clips are outputs of preprocess_audio: no filter, no trim
CLIP_OK, CLIP_EMPTY, CLIP_NONFINITE, CLIP_TOO_SHORT = 0, 1, 2, 3
ABI_VERSION = 1
ERR_INVALID_ARG, ERR_CUDA, ERR_WORKSPACE, ERR_UNSUPPORTED = 1, 2, 3, 4

EXPORTS = [
    "afx_config_default", "afx_abi_version", "afx_last_error", "afx_create", "afx_destroy",
    "afx_summary_width", "afx_frame_width", "afx_max_frames", "afx_workspace_bytes",
    "afx_workspace_min_bytes", "afx_extract", "afx_extract_host", "afx_extract_host_packed",
    "afx_launch_count", "afx_profile_enable", "afx_profile_read", "afx_preprocess", "afx_frame_extras", "afx_wav_probe_batch", "afx_wav_load_batch", "afx_f0_postprocess", "afx_resample_length", "afx_resample_filter", "afx_resample", "afx_dtw_frame_distance", "afx_dtw_accumulated_cost",
    "afx_clip_kernel_variant", "afx_debug_header_bytes", "afx_extract_static", "afx_static_workspace_capacity",
]
PHASES = ("trim", "frames", "reserved2", "reserved3", "pcm", "yin", "viterbi", "reserved7")
N_PHASES = 8
SAMPLE_F32, SAMPLE_S16 = 0, 1
EXTRAS_WIDTH = 5


class AfxConfig(C.Structure):
    _fields_ = [
        ("sr", C.c_int32), ("frame_length", C.c_int32), ("hop_length", C.c_int32), ("n_mfcc", C.c_int32),
        ("f0_min", C.c_double), ("f0_max", C.c_double),
        ("pre_emphasis", C.c_float), ("trim_top_db", C.c_float),
        ("n_mels", C.c_int32), ("mfcc_top_db", C.c_float), ("device", C.c_int32),
        ("reserved", C.c_int32 * 7),
    ]


class AfxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"afx error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    """Load libafx_b200.so once; fail loudly when it is missing (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA extension has not been built. Run "
            "`python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc); there is no CPU fallback.")
    L = C.CDLL(str(LIB_PATH))
    vp, i64, i32, u32, sz = C.c_void_p, C.c_int64, C.c_int32, C.c_uint32, C.c_size_t
    L.afx_config_default.argtypes = [C.POINTER(AfxConfig)]; L.afx_config_default.restype = None
    L.afx_abi_version.argtypes = []; L.afx_abi_version.restype = C.c_int
    L.afx_last_error.argtypes = []; L.afx_last_error.restype = C.c_char_p
    L.afx_create.argtypes = [C.POINTER(AfxConfig), C.POINTER(vp)]; L.afx_create.restype = C.c_int
    L.afx_destroy.argtypes = [vp]; L.afx_destroy.restype = C.c_int
    L.afx_summary_width.argtypes = [vp]; L.afx_summary_width.restype = C.c_int
    L.afx_frame_width.argtypes = [vp]; L.afx_frame_width.restype = C.c_int
    L.afx_max_frames.argtypes = [vp, i64]; L.afx_max_frames.restype = i64
    L.afx_workspace_bytes.argtypes = [vp, i64, vp]; L.afx_workspace_bytes.restype = sz
    L.afx_workspace_min_bytes.argtypes = [vp, i64, vp]; L.afx_workspace_min_bytes.restype = sz
    L.afx_extract.argtypes = [vp, vp, vp, vp, i64, u32, vp, vp, vp, vp, sz, vp]; L.afx_extract.restype = C.c_int
    L.afx_extract_static.argtypes = [vp, vp, vp, vp, i64, u32, vp, vp, vp]; L.afx_extract_static.restype = C.c_int
    L.afx_static_workspace_capacity.argtypes = []; L.afx_static_workspace_capacity.restype = sz
    L.afx_preprocess.argtypes = [vp, vp, vp, vp, i64, vp, vp, vp, vp, sz, vp]; L.afx_preprocess.restype = C.c_int
    L.afx_frame_extras.argtypes = [vp, vp, i64, vp, vp, vp, vp, vp]; L.afx_frame_extras.restype = C.c_int
    L.afx_wav_probe_batch.argtypes = [vp, i64, C.c_int, vp, vp, vp, vp, vp]; L.afx_wav_probe_batch.restype = C.c_int
    L.afx_wav_load_batch.argtypes = [vp, i64, C.c_int, C.c_int, vp, vp, vp, vp]; L.afx_wav_load_batch.restype = C.c_int
    L.afx_f0_postprocess.argtypes = [vp, i64, vp, vp, vp, vp, vp, vp, vp]; L.afx_f0_postprocess.restype = C.c_int
    L.afx_resample_length.argtypes = [vp, C.c_int32, i64]; L.afx_resample_length.restype = i64
    L.afx_clip_kernel_variant.argtypes = [C.c_int32, C.c_int32, C.c_int32, vp]; L.afx_clip_kernel_variant.restype = C.c_int
    L.afx_resample_filter.argtypes = [C.c_int32, C.c_int32, vp, i64, vp, vp, vp, vp]; L.afx_resample_filter.restype = i64
    L.afx_resample.argtypes = [vp, vp, vp, vp, i64, C.c_int32, vp, vp, vp]; L.afx_resample.restype = C.c_int
    L.afx_dtw_frame_distance.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_int32, i64, vp, vp, vp]; L.afx_dtw_frame_distance.restype = C.c_int
    L.afx_dtw_accumulated_cost.argtypes = [vp, vp, vp, vp, vp, i64, vp, vp]; L.afx_dtw_accumulated_cost.restype = C.c_int
    L.afx_extract_host.argtypes = [vp, vp, vp, i64, u32, vp, vp]; L.afx_extract_host.restype = C.c_int
    L.afx_extract_host_packed.argtypes = [vp, vp, C.c_int, i64, vp, vp, i64, u32, vp, vp]
    L.afx_extract_host_packed.restype = C.c_int
    L.afx_launch_count.argtypes = []; L.afx_launch_count.restype = i64
    L.afx_profile_enable.argtypes = [vp, C.c_int]; L.afx_profile_enable.restype = C.c_int
    L.afx_profile_read.argtypes = [vp, vp, vp, C.c_int]; L.afx_profile_read.restype = C.c_int
    L.afx_debug_header_bytes.argtypes = [i64]; L.afx_debug_header_bytes.restype = sz
    if L.afx_abi_version() != ABI_VERSION:
        raise ImportError(f"libafx_b200.so ABI {L.afx_abi_version()} != binding {ABI_VERSION}; rebuild")
    _lib = L
    return L


def check(code):
    if code != 0:
        raise AfxError(code, lib().afx_last_error().decode("utf-8", "replace"))
