"""ctypes binding of libpaddle_ray_b200.so (the C ABI declared in include/paddle_ray_b200.h).

There is deliberately no fallback: if the shared library is missing this module raises at
import time, and every compute call fails on a machine without a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libpaddle_ray_b200.so")

BOOL, INT32, INT64, FLOAT32, FLOAT64, FLOAT16, BFLOAT16 = 1, 7, 9, 10, 11, 15, 16
DTYPE_SIZE = {BOOL: 1, INT32: 4, INT64: 8, FLOAT32: 4, FLOAT64: 8, FLOAT16: 2, BFLOAT16: 2}
DTYPE_NAME = {BOOL: "bool", INT32: "int32", INT64: "int64", FLOAT32: "float32", FLOAT64: "float64",
              FLOAT16: "float16", BFLOAT16: "bfloat16"}
UNIQUE_ID_BYTES = 128


class PRBTensor(C.Structure):
    _fields_ = [("data", C.c_void_p), ("numel", C.c_int64), ("dtype", C.c_int32), ("device", C.c_int32)]


class TailConfig(C.Structure):
    _fields_ = [
        ("beta1", C.c_double), ("beta2", C.c_double), ("epsilon", C.c_double),
        ("coeff", C.c_float), ("use_adamw", C.c_int32), ("multi_precision", C.c_int32), ("use_amp", C.c_int32),
        ("use_clip", C.c_int32), ("clip_norm", C.c_float), ("auto_skip_clip", C.c_int32), ("use_mp_type", C.c_int32),
        ("write_back_grads", C.c_int32), ("zero_grads", C.c_int32), ("keep_beta_pow", C.c_int32),
        ("pass1_done", C.c_int32), ("pass1_partials", C.c_void_p), ("stage", C.c_int32),
        ("param_to_grad_slot", C.c_int32),
    ]


class SelectedRowsC(C.Structure):
    _fields_ = [("rows", C.c_void_p), ("nrows", C.c_int64), ("height", C.c_int64), ("value", PRBTensor)]


class GradEpilogue(C.Structure):
    _fields_ = [("loss_scale", C.POINTER(PRBTensor)), ("found_inf", C.POINTER(PRBTensor)), ("want_norm", C.c_int32),
                ("use_mp_type", C.c_int32), ("partial_base", C.c_void_p)]


class PRBError(RuntimeError):
    """Raised for any non-zero PRB_Status; mirrors the PADDLE_ENFORCE categories."""

    CATEGORY = {1: "InvalidArgument", 2: "PreconditionNotMet", 3: "OutOfRange", 4: "Unimplemented", 5: "External"}

    def __init__(self, status, message):
        self.status = status
        super().__init__("(%s) %s" % (self.CATEGORY.get(status, status), message))


if not os.path.exists(LIB_PATH):
    raise ImportError(
        "paddle_ray_b200: %s is missing. Build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(or `make -C paddle_ray_b200/csrc`). There is no CPU or pure-Python fallback." % LIB_PATH)

lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)

_T = C.POINTER(PRBTensor)
_vp = C.c_void_p
_i32, _i64, _f32, _f64 = C.c_int32, C.c_int64, C.c_float, C.c_double

# name -> argtypes; every function returns int32 status unless listed in _RESTYPES
SIGNATURES = {
    "prb_adamw": [_T] * 9 + [_f64] * 3 + [_f32, _f32, _i32, _i32, _i64, _i32, _i32] + [_T] * 6 + [_vp],
    "prb_adam": [_T] * 9 + [_f64] * 3 + [_i32, _i64, _i32, _i32] + [_T] * 6 + [_vp],
    "prb_merged_adam": [_T] * 8 + [_i64] + [_f64] * 3 + [_i32, _i32] + [_T] * 6 + [_vp],
    "prb_fused_adam": [_T] * 8 + [_i64, _T] + [_f64] * 3 + [_i32, _f32, _i32, _i32, _i32] + [_T] * 6 + [_vp],
    "prb_adam_dense_param_sparse_grad": [_T, C.POINTER(SelectedRowsC)] + [_T] * 7 + [_f64] * 3 + [_i32, _i64, _i32, _i32] +
                                        [_T] * 6 + [_vp],
    "prb_multiply_by_clip": [_T, _i64, _T, _vp],
    "prb_merge_selected_rows": [C.POINTER(SelectedRowsC), _vp, _vp, _T, _vp],
    "prb_check_finite_and_unscale": [_T, _i64, _T, _T, _T, _vp],
    "prb_update_loss_scaling": [_T, _i64, _T, _T, _T, _T, _i32, _i32, _f32, _f32, _i32, _T, _T, _T, _T, _vp],
    "prb_squared_l2_norm": [_T, _T, _vp],
    "prb_squared_l2_norm_noalloc": [_T, _T, _vp],
    "prb_clip_grad_by_global_norm": [_T, _i64, _f32, _i32, _i32, _T, _T, _vp],
    "prb_scale": [_T, _f64, _f32, _i32, _T, _vp],
    "prb_concat": [_T, _i64, _T, _vp],
    "prb_split": [_T, _T, _i64, _vp],
    "prb_cast": [_T, _T, _vp],
    "prb_multiply_by_scalar_tensor": [_T, _T, _T, _vp],
    "prb_assign_group_by_size": [_vp, _vp, _vp, _i64, _vp, _i64, _vp, _i64, _vp, _vp, _vp],
    "prb_comm_get_unique_id": [_vp],
    "prb_comm_init_rank": [_vp, _i32, _i32, _i32, C.POINTER(_vp)],
    "prb_comm_init_single": [_i32, C.POINTER(_vp)],
    "prb_comm_destroy": [_vp],
    "prb_comm_nranks": [_vp],
    "prb_comm_rank": [_vp],
    "prb_comm_stream": [_vp],
    "prb_comm_all_reduce": [_vp, _T, _T, _i32, _vp],
    "prb_comm_broadcast": [_vp, _T, _i32, _vp],
    "prb_comm_all_gather": [_vp, _T, _T, _vp],
    "prb_reducer_create": [_T, _i64, _vp, _vp, _i64, _vp, _i32, C.POINTER(_vp)],
    "prb_reducer_destroy": [_vp],
    "prb_reducer_set_sharded": [_vp, _i32, _vp],
    "prb_reducer_shard_ranges": [_vp, _i32, _i32, _vp, _vp, _vp, _i64, _vp],
    "prb_reducer_set_shard_layout": [_vp, _i32],
    "prb_reducer_allgather_params": [_vp, _T, _i64, _T, _vp],
    "prb_reducer_prepare_for_backward": [_vp],
    "prb_reducer_set_unused_vars": [_vp, _vp, _i64],
    "prb_reducer_mark_var_ready": [_vp, _i64, _vp],
    "prb_reducer_mark_vars_ready": [_vp, _vp, _i64, _vp],
    "prb_reducer_set_unpack": [_vp, _i32],
    "prb_reducer_set_grads_in_arena": [_vp, _i32],
    "prb_reducer_set_grad_epilogue": [_vp, C.POINTER(GradEpilogue), _T, C.POINTER(_i32)],
    "prb_grad_chunk_elems": [],
    "prb_reducer_set_deferred_wait": [_vp, _i32],
    "prb_reducer_wait_group": [_vp, _i64, _vp],
    "prb_reducer_num_groups": [_vp],
    "prb_reducer_group_of": [_vp, _i64, C.POINTER(_i64)],
    "prb_reducer_allreduce_backend": [_vp],
    "prb_reducer_allreduce_group": [_vp, _i64, _vp],
    "prb_reducer_comm_error": [_vp, C.POINTER(_i32)],
    "prb_reducer_set_exchange_ctas": [_vp, _i32],
    "prb_reducer_is_finalized": [_vp],
    "prb_reducer_bucket_view": [_vp, _i64, _T],
    "prb_reducer_launched_groups": [_vp, _vp, _i64, _vp],
    "prb_dp_step_tail": [C.POINTER(TailConfig), _T, _T, _T, _T, _T, _T, _T, _T, _i64, _vp, _vp, _T, _T, _T, _T, _vp],
    "prb_device_count": [C.POINTER(_i32)],
    "prb_set_device": [_i32],
    "prb_device_info": [_i32, C.POINTER(_i32), C.POINTER(_i64), C.POINTER(_i64), C.c_char_p],
    "prb_malloc": [_i64, C.POINTER(_vp)],
    "prb_free": [_vp],
    "prb_host_alloc": [_i64, C.POINTER(_vp)],
    "prb_host_free": [_vp],
    "prb_memcpy_h2d": [_vp, _vp, _i64, _vp],
    "prb_memcpy_d2h": [_vp, _vp, _i64, _vp],
    "prb_memcpy_d2d": [_vp, _vp, _i64, _vp],
    "prb_memset": [_vp, _i32, _i64, _vp],
    "prb_stream_create": [_i32, C.POINTER(_vp)],
    "prb_stream_destroy": [_vp],
    "prb_stream_synchronize": [_vp],
    "prb_device_synchronize": [],
    "prb_event_create": [C.POINTER(_vp)],
    "prb_event_destroy": [_vp],
    "prb_event_record": [_vp, _vp],
    "prb_event_synchronize": [_vp],
    "prb_event_elapsed_ms": [_vp, _vp, C.POINTER(_f32)],
    "prb_stream_wait_event": [_vp, _vp],
    "prb_flag_or": [_vp, _vp, _vp],
    "prb_spin": [_i64, _vp],
    "prb_flush_l2": [_vp],
    "prb_launch_count": [],
    "prb_kernel_timing": [_i32],
    "prb_kernel_timing_read": [_i32, C.POINTER(_f64), C.POINTER(_i64)],
    "prb_kernel_kind_name": [_i32],
    "prb_last_error_message": [],
    "prb_version": [],
}
_RESTYPES = {"prb_comm_stream": _vp, "prb_launch_count": _i64, "prb_last_error_message": C.c_char_p,
             "prb_version": C.c_char_p, "prb_kernel_kind_name": C.c_char_p}
# entry points whose int32 return value is a value, not a status
_VALUE_RETURN = {"prb_grad_chunk_elems", "prb_comm_nranks", "prb_comm_rank", "prb_reducer_is_finalized", "prb_reducer_num_groups", "prb_reducer_allreduce_backend"}

for _name, _args in SIGNATURES.items():
    _fn = getattr(lib, _name)  # AttributeError here == the .so does not export what the header declares
    _fn.argtypes = _args
    _fn.restype = _RESTYPES.get(_name, _i32)


C = C  # re-exported for callers that build ctypes arguments


def last_error() -> str:
    return lib.prb_last_error_message().decode()


def check(status: int):
    if status != 0:
        raise PRBError(status, last_error())


def call(name, *args):
    check(getattr(lib, name)(*args))


def tensor_array(tensors):
    """list of PRBTensor (or objects with .prb) -> contiguous C array (None for empty/None)."""
    if tensors is None:
        return None
    arr = (PRBTensor * max(len(tensors), 1))()
    for i, t in enumerate(tensors):
        p = t.prb if hasattr(t, "prb") else t
        arr[i].data, arr[i].numel, arr[i].dtype, arr[i].device = p.data, p.numel, p.dtype, p.device
    return arr
