"""ctypes binding of the CUDA library (libtextblaster_b200.so, built in-tree by
``__graft_entry__.build()``).  There is no CPU path: a missing library or a missing CUDA device
raises, it never falls back."""
from __future__ import annotations

import ctypes as C
import os

from . import _abi
from .error import ConfigValidationError, PipelineError, Unexpected

_HERE = os.path.dirname(os.path.abspath(__file__))
# TB_LIB_PATH: load another build of the library (A/B measurements of build-time settings, tools/build_variant.sh)
LIB_PATH = os.environ.get("TB_LIB_PATH") or os.path.join(_HERE, "libtextblaster_b200.so")
_lib = None

EXPORTS = [
    "tb_abi_version", "tb_last_error", "tb_step_name", "tb_ctx_create", "tb_ctx_destroy", "tb_pipeline_create",
    "tb_pipeline_destroy", "tb_pipeline_run_batch", "tb_pipeline_run_batch_device", "tb_batch_out_release",
    "tb_pipeline_set_option", "tb_pipeline_kernel_times", "tb_pipeline_reset_kernel_times", "tb_gather_content",
]


class NativeLibraryMissing(PipelineError):
    def __str__(self):
        return f"Unexpected error: {self.message}"


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NativeLibraryMissing(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). textblaster_b200 has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    L.tb_abi_version.restype = C.c_int
    L.tb_last_error.restype = C.c_char_p
    L.tb_last_error.argtypes = [C.c_void_p]
    L.tb_step_name.restype = C.c_char_p
    L.tb_step_name.argtypes = [C.c_int32]
    L.tb_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    L.tb_ctx_destroy.argtypes = [C.c_void_p]
    L.tb_pipeline_create.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p)]
    L.tb_pipeline_destroy.argtypes = [C.c_void_p]
    L.tb_pipeline_run_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.POINTER(_abi.tb_batch_out)]
    L.tb_pipeline_run_batch_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.POINTER(_abi.tb_batch_out)]
    L.tb_batch_out_release.argtypes = [C.c_void_p, C.POINTER(_abi.tb_batch_out)]
    L.tb_pipeline_set_option.argtypes = [C.c_void_p, C.c_int32, C.c_int64]
    L.tb_pipeline_kernel_times.restype = C.c_int
    L.tb_pipeline_kernel_times.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p), C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    L.tb_pipeline_reset_kernel_times.argtypes = [C.c_void_p]
    L.tb_gather_content.restype = C.c_int
    L.tb_gather_content.argtypes = [C.POINTER(_abi.tb_batch_out), C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int64,
                                    C.c_void_p, C.POINTER(C.c_int64), C.c_int]
    if L.tb_abi_version() != _abi.TB_ABI_VERSION:
        raise Unexpected(f"ABI mismatch: library {L.tb_abi_version()} vs binding {_abi.TB_ABI_VERSION}")
    _lib = L
    return L


def last_error() -> str:
    return lib().tb_last_error(None).decode("utf-8", "replace")


def check(rc: int):
    if rc == _abi.TB_OK:
        return
    msg = last_error()
    if rc == _abi.TB_ERR_CONFIG_VALIDATION:
        raise ConfigValidationError(msg)
    raise Unexpected(f"textblaster_b200 native error {rc}: {msg}")
