"""ctypes binding of the C ABI in ``include/rm_b200.h`` (``csrc/librm_b200.so``).

There is no fallback: if the shared library is missing it is built with nvcc, and if that is
impossible, or a call is made without a CUDA device, the call raises.  Structures mirror the
header field by field.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# RM_B200_LIB: alternative build of the same ABI (kernel tuning experiments)
LIB_PATH = os.environ.get("RM_B200_LIB") or os.path.join(_HERE, "csrc", "librm_b200.so")

POSE_STRIDE = 8
REDUCE_FLOATS = 8
MARCH_BWD_STATIC_MAX_FLOATS = 8388608      # include/rm_b200.h: RM_MARCH_BWD_STATIC_MAX_FLOATS
ABI_VERSION = 7


class RmTape(C.Structure):
    _fields_ = [("code_host", C.c_void_p), ("par", C.c_void_p), ("n_instr", C.c_int32),
                ("n_par", C.c_int32), ("max_depth", C.c_int32), ("n_nodes", C.c_int32)]


class RmCamera(C.Structure):
    _fields_ = [("n_views", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
                ("focal_length", C.c_float), ("sensor_width", C.c_float),
                ("sensor_height", C.c_float)]


class RmRenderOut(C.Structure):
    _fields_ = [("origin", C.c_void_p), ("dirs", C.c_void_p), ("pos", C.c_void_p),
                ("sdf", C.c_void_p), ("normals", C.c_void_p), ("laplacian", C.c_void_p),
                ("image", C.c_void_p * 8), ("rgba", C.c_void_p), ("reduce", C.c_void_p)]


class RmRenderArgs(C.Structure):
    _fields_ = [("ray_begin", C.c_int64), ("n_rays", C.c_int64), ("steps", C.c_int32),
                ("bounces", C.c_int32), ("normals_eps", C.c_float), ("hit_tau", C.c_float),
                ("lift_delta", C.c_float), ("mode_mask", C.c_uint32),
                ("display_mode", C.c_int32), ("degree", C.c_int32), ("early_exit", C.c_int32),
                ("fast_math", C.c_int32), ("cmap", C.c_void_p), ("bwd_converge_tol", C.c_float),
                ("strip_rays", C.c_int64), ("strip_stride", C.c_int64)]


class RmRenderGrads(C.Structure):
    _fields_ = [("origin", C.c_void_p), ("dirs", C.c_void_p), ("pos", C.c_void_p),
                ("sdf", C.c_void_p), ("normals", C.c_void_p), ("laplacian", C.c_void_p)]


class RmLossArgs(C.Structure):
    _fields_ = [("target_lambertian", C.c_void_p), ("target_depth", C.c_void_p),
                ("weight_lambertian", C.c_float), ("weight_depth", C.c_float), ("loss_out", C.c_void_p),
                ("scratch_floats", C.c_int64)]


_SIGNATURES = {
    "rm_abi_version": (C.c_int, []),
    "rm_last_error": (C.c_char_p, []),
    "rm_sdf_eval": (C.c_int, [C.POINTER(RmTape), C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
    "rm_camera_rays": (C.c_int, [C.POINTER(RmCamera), C.c_void_p, C.c_int64, C.c_int64,
                                 C.c_void_p, C.c_void_p, C.c_void_p]),
    "rm_march": (C.c_int, [C.POINTER(RmTape), C.c_void_p, C.c_void_p, C.c_int64, C.c_int32,
                           C.c_int32, C.c_void_p, C.c_void_p]),
    "rm_camera_update": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_float, C.c_float,
                                   C.c_void_p, C.c_void_p]),
    "rm_normals": (C.c_int, [C.POINTER(RmTape), C.c_void_p, C.c_int64, C.c_float, C.c_void_p,
                             C.c_void_p, C.c_void_p]),
    "rm_render": (C.c_int, [C.POINTER(RmTape), C.POINTER(RmCamera), C.c_void_p,
                            C.POINTER(RmRenderArgs), C.POINTER(RmRenderOut), C.c_void_p]),
    "rm_reduce_init": (C.c_int, [C.c_void_p, C.c_void_p]),
    "rm_shade_finalize": (C.c_int, [C.c_void_p, C.c_uint32, C.c_int64, C.POINTER(C.c_void_p),
                                    C.c_void_p, C.c_int32, C.c_void_p]),
    "rm_planes_to_f16": (C.c_int, [C.c_uint32, C.c_int64, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                   C.c_void_p]),
    "rm_sdf_vjp": (C.c_int, [C.POINTER(RmTape), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                             C.c_void_p, C.c_void_p]),
    "rm_march_bwd": (C.c_int, [C.POINTER(RmTape), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                               C.c_int32, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                               C.c_void_p]),
    "rm_march_bwd_static": (C.c_int, [C.POINTER(RmTape), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                      C.c_int32, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "rm_bwd_scratch_floats": (C.c_int64, [C.c_int64, C.c_int32]),
    "rm_render_bwd": (C.c_int, [C.POINTER(RmTape), C.POINTER(RmCamera), C.c_void_p,
                                C.POINTER(RmRenderArgs), C.POINTER(RmRenderGrads), C.c_void_p,
                                C.c_void_p, C.c_void_p, C.c_void_p]),
    "rm_render_loss_bwd": (C.c_int, [C.POINTER(RmTape), C.POINTER(RmCamera), C.c_void_p,
                                     C.POINTER(RmRenderArgs), C.POINTER(RmLossArgs), C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p]),
    "rm_loss_scratch_floats": (C.c_int64, [C.POINTER(RmTape), C.c_int64, C.c_int32]),
    "rm_probe_ffma": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "rm_probe_issue": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "rm_probe_math": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]),
}

_lib = None


def exported_symbols():
    """Names declared in include/rm_b200.h that the library must export (used by tests)."""
    return list(_SIGNATURES.keys())


def lib() -> C.CDLL:
    """Loads (building first if needed) the shared library; raises if that is impossible."""
    global _lib
    if _lib is None:
        if not os.environ.get("RM_B200_LIB"):
            # (re)build when the library is missing or its sources changed (content digest kept
            # beside the .so); an alternative build named by RM_B200_LIB is taken as is
            from .csrc.build import build
            build()
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        if handle.rm_abi_version() != ABI_VERSION:
            raise RuntimeError(f"{LIB_PATH}: ABI version {handle.rm_abi_version()} != {ABI_VERSION}; rebuild")
        _lib = handle
    return _lib


class RmError(RuntimeError):
    pass


def check(status: int, what: str) -> None:
    if status != 0:
        msg = lib().rm_last_error().decode(errors="replace")
        raise RmError(f"{what} failed with status {status}: {msg}")


def require_cuda(*tensors: torch.Tensor) -> torch.device:
    """The path has no CPU implementation: every tensor must live on one CUDA device."""
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RmError(
                "ray_marching_b200 runs only on CUDA tensors (there is no CPU fallback); got a "
                f"tensor on {t.device}")
        if t.dtype != torch.float32:
            raise RmError(f"ray_marching_b200 computes in float32; got {t.dtype}")
        dev = dev or t.device
        if t.device != dev:
            raise RmError(f"tensors on different devices: {t.device} vs {dev}")
    if dev is None:
        raise RmError("no CUDA tensor given")
    return dev


def stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()
