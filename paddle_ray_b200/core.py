"""Device tensors, streams and events for the torch-free host shim.

A ``Tensor`` is the minimum the step tail needs from ``paddle.Tensor``: a typed 1-D device (or
host) buffer with ``numel``/``dtype``/``place``, numpy round-trips, and a ``.grad`` slot.  16-bit
floats travel through numpy as ``uint16`` bit patterns (numpy has no bfloat16).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import BFLOAT16, BOOL, FLOAT16, FLOAT32, FLOAT64, INT32, INT64, PRBTensor, call, lib

_DTYPE_ALIASES = {
    "float32": FLOAT32, "float64": FLOAT64, "float16": FLOAT16, "bfloat16": BFLOAT16, "uint16": BFLOAT16,
    "bool": BOOL, "int32": INT32, "int64": INT64,
}
_NP_STORAGE = {FLOAT32: np.float32, FLOAT64: np.float64, FLOAT16: np.uint16, BFLOAT16: np.uint16, BOOL: np.bool_,
               INT32: np.int32, INT64: np.int64}


def dtype_code(dtype) -> int:
    if isinstance(dtype, int):
        return dtype
    if isinstance(dtype, str):
        return _DTYPE_ALIASES[dtype]
    return _DTYPE_ALIASES[np.dtype(dtype).name]


def device_count() -> int:
    n = C.c_int32()
    call("prb_device_count", C.byref(n))
    return n.value


_current_device = 0


def set_device(device: int):
    global _current_device
    call("prb_set_device", int(device))
    _current_device = int(device)


def current_device() -> int:
    return _current_device


def device_info(device: int | None = None) -> dict:
    sm, l2, mem = C.c_int32(), C.c_int64(), C.c_int64()
    name = C.create_string_buffer(128)
    call("prb_device_info", current_device() if device is None else device, C.byref(sm), C.byref(l2), C.byref(mem), name)
    return {"sm_count": sm.value, "l2_bytes": l2.value, "total_mem": mem.value, "name": name.value.decode()}


class Stream:
    def __init__(self, high_priority: bool = False, handle=None):
        self._owned = handle is None
        if handle is None:
            h = C.c_void_p()
            call("prb_stream_create", int(high_priority), C.byref(h))
            handle = h.value
        self.handle = handle

    def synchronize(self):
        call("prb_stream_synchronize", self.handle)

    def wait_event(self, event: "Event"):
        call("prb_stream_wait_event", self.handle, event.handle)

    def __del__(self):
        try:
            if self._owned and self.handle:
                lib.prb_stream_destroy(self.handle)
        except Exception:
            pass


class Event:
    def __init__(self):
        h = C.c_void_p()
        call("prb_event_create", C.byref(h))
        self.handle = h.value

    def record(self, stream: Stream | None = None):
        call("prb_event_record", self.handle, _stream_handle(stream))

    def synchronize(self):
        call("prb_event_synchronize", self.handle)

    def elapsed_ms(self, end: "Event") -> float:
        ms = C.c_float()
        call("prb_event_elapsed_ms", self.handle, end.handle, C.byref(ms))
        return ms.value

    def __del__(self):
        try:
            if self.handle:
                lib.prb_event_destroy(self.handle)
        except Exception:
            pass


def synchronize():
    call("prb_device_synchronize")


_current_streams: dict = {}


def current_stream() -> Stream:
    """The calc stream of the current device (created on first use; non-blocking).

    Every op of the shim and every Tensor copy defaults to it, so default-argument code is
    stream-ordered end to end without ever touching the legacy NULL stream.
    """
    dev = current_device()
    s = _current_streams.get(dev)
    if s is None:
        s = _current_streams[dev] = Stream()
    return s


def set_stream(stream: Stream) -> Stream:
    prev = current_stream()
    _current_streams[current_device()] = stream
    return prev


def _stream_handle(stream):
    if stream is None:
        return current_stream().handle
    return stream.handle if isinstance(stream, Stream) else stream


class Tensor:
    """1-D typed buffer on the device (``device >= 0``) or on the host (``device == -1``)."""

    def __init__(self, numel: int, dtype, device: int | None = None, *, data_ptr=None, owner=None, name=None,
                 shape=None):
        self.dtype = dtype_code(dtype)
        self.numel = int(numel)
        self.device = current_device() if device is None else int(device)
        self.name = name
        self.shape = list(shape) if shape is not None else [self.numel]
        self.grad: Tensor | None = None
        self.need_clip = True
        self.stop_gradient = False
        self.optimize_attr = {"learning_rate": 1.0}
        self.is_sparse_grad = False
        self._reduced_grad = None      # view of this parameter's slot in a DataParallel bucket arena
        self._bucket = (None, 0)       # (DataParallel, launch-order group index) when gradients are consumed there
        self._owner = owner  # keeps the backing allocation alive for views
        self._host = None
        self._owned = False
        if data_ptr is not None:
            self.data_ptr = int(data_ptr)
        elif self.device < 0:
            self._host = np.zeros(self.numel, dtype=_NP_STORAGE[self.dtype])
            self.data_ptr = self._host.ctypes.data
        else:
            p = C.c_void_p()
            call("prb_malloc", self.nbytes, C.byref(p))
            self.data_ptr = p.value
            self._owned = True

    # ---- construction
    @classmethod
    def from_numpy(cls, arr: np.ndarray, dtype=None, device: int | None = None, stream=None, name=None) -> "Tensor":
        """``arr`` holds values for float32/int/bool dtypes, or uint16 bit patterns for 16-bit floats."""
        arr = np.ascontiguousarray(arr)
        code = dtype_code(dtype) if dtype is not None else dtype_code(arr.dtype)
        assert arr.dtype == _NP_STORAGE[code], "storage dtype %s does not match %s" % (arr.dtype, _lib.DTYPE_NAME[code])
        t = cls(arr.size, code, device, name=name, shape=arr.shape)
        t.copy_from_numpy(arr, stream)
        return t

    @classmethod
    def zeros(cls, numel: int, dtype, device: int | None = None, stream=None, name=None) -> "Tensor":
        t = cls(numel, dtype, device, name=name)
        t.zero_(stream)
        return t

    @classmethod
    def full(cls, numel: int, value, dtype, device: int | None = None, name=None) -> "Tensor":
        code = dtype_code(dtype)
        return cls.from_numpy(np.full(numel, value, dtype=_NP_STORAGE[code]), code, device, name=name)

    # ---- properties
    @property
    def nbytes(self) -> int:
        return self.numel * _lib.DTYPE_SIZE[self.dtype]

    @property
    def prb(self) -> PRBTensor:
        return PRBTensor(self.data_ptr, self.numel, self.dtype, self.device)

    @property
    def place(self) -> str:
        return "cpu" if self.device < 0 else "gpu:%d" % self.device

    def view(self, offset_elems: int, numel: int) -> "Tensor":
        return Tensor(numel, self.dtype, self.device, owner=self,
                      data_ptr=self.data_ptr + offset_elems * _lib.DTYPE_SIZE[self.dtype])

    # ---- data movement
    def copy_from_numpy(self, arr: np.ndarray, stream=None):
        arr = np.ascontiguousarray(arr).reshape(-1)
        assert arr.size == self.numel and arr.dtype == _NP_STORAGE[self.dtype]
        if self.device < 0:
            self._host[:] = arr
            return
        call("prb_memcpy_h2d", self.data_ptr, arr.ctypes.data, self.nbytes, _stream_handle(stream))
        call("prb_stream_synchronize", _stream_handle(stream))  # `arr` may be a temporary

    def numpy(self, stream=None) -> np.ndarray:
        out = np.empty(self.numel, dtype=_NP_STORAGE[self.dtype])
        if self.device < 0:
            C.memmove(out.ctypes.data, self.data_ptr, self.nbytes)
            return out
        call("prb_memcpy_d2h", out.ctypes.data, self.data_ptr, self.nbytes, _stream_handle(stream))
        call("prb_stream_synchronize", _stream_handle(stream))
        return out

    def item(self, stream=None):
        return self.numpy(stream)[0]

    def zero_(self, stream=None):
        if self.device < 0:
            C.memset(self.data_ptr, 0, self.nbytes)
        else:
            call("prb_memset", self.data_ptr, 0, self.nbytes, _stream_handle(stream))
        return self

    def copy_(self, other: "Tensor", stream=None):
        assert other.numel == self.numel and other.dtype == self.dtype
        fn = {(False, False): "prb_memcpy_d2d", (True, False): "prb_memcpy_h2d", (False, True): "prb_memcpy_d2h"}
        key = (other.device < 0, self.device < 0)
        if key == (True, True):
            C.memmove(self.data_ptr, other.data_ptr, self.nbytes)
        else:
            call(fn[key], self.data_ptr, other.data_ptr, self.nbytes, _stream_handle(stream))
        return self

    def astype(self, dtype, stream=None) -> "Tensor":
        code = dtype_code(dtype)
        out = Tensor(self.numel, code, self.device)
        a, b = self.prb, out.prb
        call("prb_cast", C.byref(a), C.byref(b), _stream_handle(stream))
        return out

    def clear_gradient(self, set_to_zero: bool = True, stream=None):
        if isinstance(self.grad, SelectedRows):  # a sparse gradient is produced anew by every backward
            self.grad = None
        elif self.grad is not None:
            if set_to_zero:
                self.grad.zero_(stream)
            else:
                self.grad = None

    def __del__(self):
        try:
            if self._owned and self.data_ptr:
                lib.prb_free(self.data_ptr)
        except Exception:
            pass

    def __repr__(self):
        return "Tensor(numel=%d, dtype=%s, place=%s)" % (self.numel, _lib.DTYPE_NAME[self.dtype], self.place)


class SelectedRows:
    """A sparse gradient (ref paddle/phi/core/selected_rows.h): ``rows`` -- a host int64 array of row indices into a
    [height, width] parameter, any order, duplicates allowed -- and ``value``, the dense [len(rows), width] device
    tensor.  The embedding-table case: only the looked-up rows carry a gradient."""

    def __init__(self, rows, height: int, value: Tensor):
        self.rows = np.ascontiguousarray(rows, np.int64).reshape(-1)
        self.height = int(height)
        self.value = value
        if self.rows.size:
            assert value.numel % self.rows.size == 0, "value must hold len(rows) x width elements"

    @property
    def width(self) -> int:
        return self.value.numel // self.rows.size if self.rows.size else 0

    @property
    def dtype(self):
        return self.value.dtype

    @property
    def device(self):
        return self.value.device

    def is_selected_rows(self) -> bool:
        return True

    def c_struct(self):
        """-> (SelectedRowsC, keepalive) for the C ABI."""
        s = _lib.SelectedRowsC(self.rows.ctypes.data, self.rows.size, self.height, self.value.prb)
        return s, (self.rows, self.value)

    def __repr__(self):
        return "SelectedRows(rows=%d, height=%d, width=%d, dtype=%s)" % (
            self.rows.size, self.height, self.width, _lib.DTYPE_NAME[self.dtype])


class PinnedBuffer:
    """Page-locked host memory exposed as a numpy array (for the timed H2D/D2H legs)."""

    def __init__(self, numel: int, np_dtype):
        self.nbytes = int(numel) * np.dtype(np_dtype).itemsize
        p = C.c_void_p()
        call("prb_host_alloc", self.nbytes, C.byref(p))
        self.ptr = p.value
        buf = (C.c_char * self.nbytes).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=np_dtype, count=numel)

    def __del__(self):
        try:
            self.array = None
            lib.prb_host_free(self.ptr)
        except Exception:
            pass
