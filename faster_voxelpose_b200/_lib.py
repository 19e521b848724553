"""ctypes binding of libfvp_b200.so (the C ABI declared in include/fvp_b200.h).

The product path has no CPU fallback: if the library is missing or a call fails this module
raises.  Tensors cross the boundary as ``data_ptr()`` integers plus sizes; the current torch CUDA
stream is passed as the ``stream`` argument of every call.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FVP_B200_LIB") or os.path.join(_HERE, "libfvp_b200.so")   # override: kernel A/B builds
ABI_VERSION = 13
CAM_STRIDE = 24
WIN_STRIDE = 12

MODE_DIRECT = 0
FLAG_CACHED_GRID = 16
FLAG_PLANES_NHWC = 32
FLAG_TAP_TABLE = 64
FLAG_SPLIT17 = 128
TAP_ENTRY_BYTES = 20
OPT_WALK_VARIANT = 0
NMS_STATIC_MAX_CELLS = 1 << 20  # FVP_NMS_STATIC_MAX_CELLS: workspace of fvp_nms_topk_static, floats

_vp = C.c_void_p
_i = C.c_int


class Geom(C.Structure):
    """Mirror of ``struct fvp_geom``."""
    _fields_ = [("V", C.c_int32), ("J", C.c_int32), ("H", C.c_int32), ("W", C.c_int32),
                ("clamp_max", C.c_float), ("t", C.c_float * 6), ("hm_w", C.c_float), ("hm_h", C.c_float),
                ("img_w", C.c_float), ("img_h", C.c_float), ("wm1", C.c_float), ("hm1", C.c_float)]


_F3 = C.c_float * 3
_I3 = C.c_int32 * 3

_SIGNATURES = {
    "fvp_abi_version": ([], _i),
    "fvp_last_error": ([], C.c_char_p),
    "fvp_device_sm_count": ([C.POINTER(_i)], _i),
    "fvp_set_option": ([_i, _i], _i),
    "fvp_get_option": ([_i, C.POINTER(_i)], _i),
    "fvp_sample_grid": ([_vp, C.POINTER(Geom), _vp, _vp, _vp, _i, _i, _i, _vp, _vp], _i),
    "fvp_tap_table": ([_vp, C.POINTER(Geom), _vp, _vp, _vp, _i, _i, _i, _vp, _vp], _i),
    "fvp_heatmaps_channels_last": ([_vp, _i, _i, _i, _i, _i, _vp, _vp], _i),
    "fvp_heatmaps_rowpair": ([_vp, _i, _i, _i, _i, _i, _vp, _vp], _i),
    "fvp_occupancy_words": ([_i, _i, _i], C.c_long),
    "fvp_heatmaps_rowpair_occ": ([_vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp], _i),
    "fvp_individual_windows": ([_vp, _vp, _vp, _i, _F3, _F3, _I3, _F3, _F3, _I3, _vp, _vp, _vp], _i),
    "fvp_project_cubes": ([_vp, _i, _i, _vp, _vp, C.POINTER(Geom), _vp, _vp, _vp, _i, _i, _i, _vp, _i, _vp,
                           _i, _i, _i, _vp, C.c_uint, _vp], _i),
    "fvp_project_cubes_occ": ([_vp, _i, _i, _vp, _vp, C.POINTER(Geom), _vp, _vp, _vp, _i, _i, _i, _vp, _i, _vp,
                               _i, _i, _i, _vp, C.c_uint, _vp, _vp], _i),
    "fvp_project_planes": ([_vp, _i, _i, _vp, _vp, C.POINTER(Geom), _vp, _vp, _vp, _i, _i, _i, _vp, _i, _vp,
                            _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, C.c_uint, _vp], _i),
    "fvp_project_planes_occ": ([_vp, _i, _i, _vp, _vp, C.POINTER(Geom), _vp, _vp, _vp, _i, _i, _i, _vp, _i, _vp,
                                _i, _i, _i, _i, _i, _vp, _vp, _vp, _vp, C.c_uint, _vp, _vp], _i),
    "fvp_heatmaps_rowpair_split17": ([_vp, _i, _i, _i, _i, _vp, _vp, _vp], _i),
    "fvp_topk_windows": ([_vp, _i, _i, _i, _i, _i, _vp, _vp], _i),
    "fvp_planes3": ([_vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp], _i),
    "fvp_planes_combine_peers": ([C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), _i, C.POINTER(C.c_int32), C.c_long,
                                  C.c_long, C.c_long, _vp, _vp, _vp, _vp], _i),
    "fvp_planes_combine_multimem": ([_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, C.POINTER(C.c_int32), C.c_long, C.c_long,
                                     C.c_long, _vp], _i),
    "fvp_nms_topk": ([_vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp], _i),
    "fvp_nms_topk_static": ([_vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp], _i),
    "fvp_gather_topk": ([_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp], _i),
    "fvp_proposal_records": ([_vp, _vp, _vp, _vp, _i, _i, _F3, _F3, C.c_float, _vp, _vp], _i),
    "fvp_softargmax": ([_vp, _vp, _i, _i, _i, C.c_float, _vp, _vp, _vp], _i),
    "fvp_weightnet_features": ([_vp, _i, _i, _vp, _vp, _i, _vp, _vp], _i),
    "fvp_weightnet_features_tf32": ([_vp, _i, _i, _vp, _vp, _i, _vp, _vp], _i),
    "fvp_pose_match": ([_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp], _i),
    "fvp_splat_heatmaps": ([_vp, _vp, _vp, _i, _i, _i, C.c_float, _i, _i, _i, _i, _vp, _vp], _i),
    "fvp_c2c_forward": ([_vp, _i, _i, _i, _vp, _vp, _i, _vp, _vp], _i),
    "fvp_res2d_supported": ([_i, _i, _i], _i),
    "fvp_res2d_profile": ([_vp], _i),
    "fvp_res2d_block_tf32": ([_vp, _i, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp], _i),
    "fvp_conv7x7_c16_tf32": ([_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp], _i),
    "fvp_conv7x7_tc_supported": ([_i, _i, _i], _i),
    "fvp_conv7x7_tc_tf32": ([_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp], _i),
    "fvp_res_tail_supported": ([_i, _i, _i], _i),
    "fvp_res_tail_profile": ([_vp], _i),
    "fvp_res_tail_tf32": ([_vp, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _vp, _vp], _i),
    "fvp_conv7x7_c20_tf32": ([_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp], _i),
    "fvp_maxpool2_nhwc": ([_vp, _i, _i, _i, _i, _vp, _vp], _i),
    "fvp_upshuffle_add": ([_vp, _vp, _i, _i, _i, _i, _vp, _vp], _i),
    "fvp_pointwise_planar": ([_vp, _vp, _i, _i, _i, _i, _vp, _vp, _i, _vp, _vp], _i),
    "fvp_pointwise_planar_res": ([_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _i, _vp, _vp], _i),
    "fvp_fuse": ([_vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp], _i),
}
EXPORTS = tuple(_SIGNATURES)

_lib = None


class FvpError(RuntimeError):
    pass


def load() -> C.CDLL:
    """Load the shared library (raises if it has not been built -- there is no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FvpError(f"{LIB_PATH} is missing: build it with `python -m faster_voxelpose_b200.build` "
                       "(there is no CPU or PyTorch fallback for the projection path)")
    lib = C.CDLL(LIB_PATH)
    for name, (argtypes, restype) in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = restype
    if lib.fvp_abi_version() != ABI_VERSION:
        raise FvpError(f"libfvp_b200.so ABI {lib.fvp_abi_version()} != expected {ABI_VERSION}; rebuild")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise FvpError(f"{what} failed ({rc}): {load().fvp_last_error().decode(errors='replace')}")


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def stream_handle(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(t: torch.Tensor, name: str, dtype=torch.float32, contiguous: bool = True) -> torch.Tensor:
    if not t.is_cuda:
        raise FvpError(f"{name} must live on a CUDA device (got {t.device}); this package has no CPU path")
    if t.dtype != dtype:
        raise FvpError(f"{name} must be {dtype} (got {t.dtype})")
    return t.contiguous() if contiguous else t


def f3(vals) -> "_F3":
    return _F3(*[float(v) for v in vals])


def i3(vals) -> "_I3":
    return _I3(*[int(v) for v in vals])


def set_option(key: int, value: int) -> None:
    check(load().fvp_set_option(int(key), int(value)), "fvp_set_option")


def get_option(key: int) -> int:
    out = _i(0)
    check(load().fvp_get_option(int(key), C.byref(out)), "fvp_get_option")
    return out.value


def sm_count() -> int:
    out = _i(0)
    check(load().fvp_device_sm_count(C.byref(out)), "fvp_device_sm_count")
    return out.value
