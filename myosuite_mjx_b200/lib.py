"""ctypes binding of libmyostep.so (include/myostep.h).  Fails loudly: there is no CPU or eager fallback."""
from __future__ import annotations

import ctypes
import os
import subprocess

from . import desc as _desc

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MYO_LIB") or os.path.join(_HERE, "libmyostep.so")   # MYO_LIB: developer knob (A/B builds)


class MyoError(RuntimeError):
    pass


def build(force=False):
    """Compile the CUDA extension in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    src = os.path.join(_HERE, "csrc")
    cmd = ["make", "-C", src, "-j4"] + (["-B"] if force else [])
    subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise MyoError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
        L = ctypes.CDLL(LIB_PATH)
        vp, i32, i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64
        L.myo_last_error.restype = ctypes.c_char_p
        L.myo_launch_count.restype = i64
        L.myo_field_name.restype = ctypes.c_char_p
        L.myo_field_name.argtypes = [i32]
        L.myo_field_is_int.argtypes = [i32]
        L.myo_field_itemsize.argtypes = [i32]
        L.myo_model_create.argtypes = [ctypes.POINTER(_desc.ModelDesc), i32, ctypes.POINTER(vp)]
        L.myo_model_destroy.argtypes = [vp]
        L.myo_model_destroy.restype = None
        L.myo_batch_create.argtypes = [vp, i32, i32, i32, ctypes.POINTER(vp)]
        L.myo_kinematics_noalloc.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp, vp]   # allocation-free forward kinematics (include/myostep.h)
        L.myo_batch_destroy.argtypes = [vp]
        L.myo_batch_destroy.restype = None
        L.myo_batch_nconmax.argtypes = [vp]
        L.myo_batch_njmax.argtypes = [vp]
        L.myo_batch_smem_bytes.argtypes = [vp]
        L.myo_batch_team.argtypes = [vp]
        L.myo_batch_ctas_per_sm.argtypes = [vp]
        L.myo_batch_engine.argtypes = [vp]
        L.myo_batch_set_pipelined.argtypes = [vp, i32]
        L.myo_batch_set_count_aware.argtypes = [vp, i32]
        L.myo_outputs_join.argtypes = [vp, vp]
        L.myo_field_count.restype = i64
        L.myo_field_count.argtypes = [vp, i32]
        L.myo_batch_bind.argtypes = [vp, i32, vp, i64]
        L.myo_forward.argtypes = [vp, vp]
        L.myo_step.argtypes = [vp, i32, vp]
        L.myo_forward_masked.argtypes = [vp, vp, vp]
        L.myo_forward_posvel.argtypes = [vp, vp, vp]
        L.myo_reset.argtypes = [vp, vp, vp]
        L.myo_step_host.argtypes = [vp, i32] + [vp] * 8
        L.myo_kat_quat.argtypes = [i32] + [vp] * 6
        _lib = L
    return _lib


def has_xla_ffi():
    """True when the library was built with the jax.ffi targets (csrc/xla_ffi.cc; needs jaxlib's XLA FFI headers,
    `make XLA_INC=...`).  This image has no jax, so the default build exports neither MyoStep nor MyoForward."""
    return hasattr(lib(), "MyoStep") and hasattr(lib(), "MyoForward")


def check(rc):
    if rc != 0:
        raise MyoError(f"libmyostep error {rc}: {lib().myo_last_error().decode()}")


EXPORTS = ["myo_model_create", "myo_model_destroy", "myo_batch_create", "myo_batch_destroy", "myo_batch_nconmax",
           "myo_batch_njmax", "myo_batch_smem_bytes", "myo_batch_team", "myo_batch_ctas_per_sm", "myo_batch_engine", "myo_batch_set_pipelined", "myo_outputs_join", "myo_field_count", "myo_field_itemsize", "myo_field_is_int", "myo_field_name", "myo_batch_bind",
           "myo_forward", "myo_forward_masked", "myo_forward_posvel", "myo_step", "myo_reset", "myo_step_host", "myo_launch_count", "myo_last_error", "myo_kat_quat",
           "myo_env_create", "myo_env_destroy", "myo_env_field_count", "myo_env_bind", "myo_env_seed", "myo_env_reset", "myo_env_step",
           "myo_env_forward", "myo_kinematics_noalloc", "myo_batch_set_count_aware"]
