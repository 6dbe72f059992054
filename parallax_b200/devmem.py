"""Device memory, streams and events WITHOUT torch: ctypes over the library's own thin wrappers of
the CUDA runtime (`cotix_device_malloc`, `cotix_memcpy`, `cotix_stream_*`, `cotix_event_*`;
include/cotix_b200.h).  This is what lets the host-side mirror run in a torch-free process (a JAX
program, a plain numpy script): `parallax_b200.lunar_lander.reset(keys, device="native")` returns a
`LunarState` whose `.data` is a `DeviceArray`, and every entry point accepts it where it accepts a
torch tensor.

A `DeviceArray` is a dense row-major array on one GPU: shape, numpy dtype, element strides (the
last dimension is contiguous; 2-D arrays may have a row stride larger than their width, i.e. be a
column slice of a wider array).  No arithmetic: the library's kernels are the arithmetic.
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np

from . import _lib

_tls = threading.local()


def device_count() -> int:
    n = C.c_int(0)
    _lib.check(_lib.lib().cotix_device_count(C.byref(n)))
    return n.value


def set_device(device: int) -> None:
    _lib.check(_lib.lib().cotix_set_device(int(device)))


class Stream:
    """A non-blocking CUDA stream owned through the C-ABI.  `with stream:` makes it the current
    stream of this thread for the native backend."""

    def __init__(self, device: int = 0):
        self.device = int(device)
        set_device(self.device)
        h = C.c_void_p()
        _lib.check(_lib.lib().cotix_stream_create(C.byref(h)))
        self.ptr = h.value or 0

    def synchronize(self) -> None:
        _lib.check(_lib.lib().cotix_stream_synchronize(self.ptr))

    def __enter__(self):
        self._prev = getattr(_tls, "stream", 0)
        _tls.stream = self.ptr
        return self

    def __exit__(self, *exc):
        _tls.stream = self._prev
        return False

    def close(self):
        if self.ptr:
            _lib.lib().cotix_stream_destroy(self.ptr)
            self.ptr = 0

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def current_stream() -> int:
    """cudaStream_t (as an int) the native backend enqueues on in this thread; 0 = default stream."""
    return getattr(_tls, "stream", 0) or 0


def synchronize() -> None:
    _lib.check(_lib.lib().cotix_stream_synchronize(current_stream()))


class Event:
    def __init__(self):
        h = C.c_void_p()
        _lib.check(_lib.lib().cotix_event_create(C.byref(h)))
        self.ptr = h.value

    def record(self, stream: int | None = None) -> None:
        _lib.check(_lib.lib().cotix_event_record(self.ptr, current_stream() if stream is None else stream))

    def elapsed_ms(self, end: "Event") -> float:
        """Milliseconds from this event to `end` (waits for `end`)."""
        ms = C.c_float(0.0)
        _lib.check(_lib.lib().cotix_event_elapsed_ms(self.ptr, end.ptr, C.byref(ms)))
        return float(ms.value)

    def __del__(self):
        try:
            if self.ptr:
                _lib.lib().cotix_event_destroy(self.ptr)
                self.ptr = None
        except Exception:
            pass


class _Block:
    """One cudaMalloc allocation; freed when the last DeviceArray view of it goes away."""

    def __init__(self, nbytes: int, device: int):
        self.device, self.nbytes = device, int(nbytes)
        set_device(device)
        h = C.c_void_p()
        _lib.check(_lib.lib().cotix_device_malloc(max(self.nbytes, 1), C.byref(h)))
        self.ptr = h.value

    def __del__(self):
        try:
            if self.ptr:
                _lib.lib().cotix_device_free(self.ptr)
                self.ptr = None
        except Exception:
            pass


class PinnedArray:
    """Page-locked host memory as a numpy array (`.array`), for asynchronous copies."""

    def __init__(self, shape, dtype=np.float32):
        self.dtype = np.dtype(dtype)
        self.shape = tuple(int(s) for s in (shape if hasattr(shape, "__len__") else (shape,)))
        n = int(np.prod(self.shape, dtype=np.int64)) * self.dtype.itemsize
        h = C.c_void_p()
        _lib.check(_lib.lib().cotix_host_malloc(max(n, 1), C.byref(h)))
        self.ptr = h.value
        buf = (C.c_char * max(n, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape, dtype=np.int64))).reshape(self.shape)

    def __del__(self):
        try:
            if self.ptr:
                self.array = None
                _lib.lib().cotix_host_free(self.ptr)
                self.ptr = None
        except Exception:
            pass


def _contig_strides(shape):
    st, acc = [], 1
    for s in reversed(shape):
        st.append(acc)
        acc *= int(s)
    return tuple(reversed(st))


class DeviceArray:
    """Dense array on the GPU (see the module docstring)."""

    is_native = True

    def __init__(self, shape, dtype=np.float32, device: int = 0, strides=None, _block=None, _offset=0):
        self.shape = tuple(int(s) for s in shape)
        self.dtype = np.dtype(dtype)
        self.strides = tuple(int(s) for s in (strides if strides is not None else _contig_strides(self.shape)))
        assert len(self.strides) == len(self.shape)
        self.device = int(device)
        if _block is None:
            span = 1 + sum((s - 1) * st for s, st in zip(self.shape, self.strides)) if self.numel() else 0
            _block = _Block(span * self.dtype.itemsize, self.device)
        self._block = _block
        self._offset = int(_offset)  # bytes

    # ---- construction ------------------------------------------------------------------
    @staticmethod
    def empty(shape, dtype=np.float32, device: int = 0, row_stride: int | None = None) -> "DeviceArray":
        shape = tuple(int(s) for s in (shape if hasattr(shape, "__len__") else (shape,)))
        strides = None
        if row_stride is not None and len(shape) >= 2 and row_stride != shape[-1]:
            st = [1, int(row_stride)]
            for s in reversed(shape[1:-1]):
                st.append(st[-1] * int(s))
            strides = tuple(reversed(st))
        return DeviceArray(shape, dtype, device, strides)

    @staticmethod
    def zeros(shape, dtype=np.float32, device: int = 0, row_stride: int | None = None) -> "DeviceArray":
        return DeviceArray.empty(shape, dtype, device, row_stride).zero_()

    @staticmethod
    def from_numpy(a: np.ndarray, device: int = 0) -> "DeviceArray":
        a = np.ascontiguousarray(a)
        out = DeviceArray(a.shape, a.dtype, device)
        out.copy_from(a)
        return out

    # ---- introspection (the subset of torch.Tensor's interface the mirror uses) -----------
    def data_ptr(self) -> int:
        return (self._block.ptr or 0) + self._offset

    def numel(self) -> int:
        return int(np.prod(self.shape, dtype=np.int64))

    def dim(self) -> int:
        return len(self.shape)

    def stride(self, i: int) -> int:
        return self.strides[i]

    def is_contiguous(self) -> bool:
        return self.strides == _contig_strides(self.shape)

    @property
    def nbytes(self) -> int:
        return self.numel() * self.dtype.itemsize

    # ---- views ---------------------------------------------------------------------------
    def __getitem__(self, idx) -> "DeviceArray":
        if not isinstance(idx, tuple):
            idx = (idx,)
        shape, strides, off = [], [], self._offset
        for d, (n, st) in enumerate(zip(self.shape, self.strides)):
            ix = idx[d] if d < len(idx) else slice(None)
            if isinstance(ix, (int, np.integer)):
                ix = int(ix) + (n if ix < 0 else 0)
                if not 0 <= ix < n:
                    raise IndexError(idx)
                off += ix * st * self.dtype.itemsize
            else:
                a, b, step = ix.indices(n)
                if step != 1:
                    raise IndexError("DeviceArray views have unit step")
                off += a * st * self.dtype.itemsize
                shape.append(max(0, b - a))
                strides.append(st)
        return DeviceArray(shape, self.dtype, self.device, strides, self._block, off)

    def view(self, dtype) -> "DeviceArray":
        dtype = np.dtype(dtype)
        assert dtype.itemsize == self.dtype.itemsize
        return DeviceArray(self.shape, dtype, self.device, self.strides, self._block, self._offset)

    def reshape(self, *shape) -> "DeviceArray":
        shape = shape[0] if len(shape) == 1 and hasattr(shape[0], "__len__") else shape
        assert self.is_contiguous(), "reshape needs a contiguous DeviceArray"
        shape = list(shape)
        if -1 in shape:
            k = shape.index(-1)
            shape[k] = self.numel() // max(1, -int(np.prod(shape)))
        assert int(np.prod(shape, dtype=np.int64)) == self.numel()
        return DeviceArray(shape, self.dtype, self.device, None, self._block, self._offset)

    # ---- data movement ---------------------------------------------------------------------
    def _rows(self):
        """(row count, row bytes, pitch bytes) of this array seen as rows of contiguous bytes."""
        if self.is_contiguous() or self.numel() == 0:
            return 1, self.nbytes, self.nbytes
        inner = _contig_strides(self.shape[1:]) if len(self.shape) > 1 else ()
        if len(self.shape) == 2 and self.strides[1] == 1:
            return self.shape[0], self.shape[1] * self.dtype.itemsize, self.strides[0] * self.dtype.itemsize
        if self.strides[1:] == inner:
            return self.shape[0], int(np.prod(self.shape[1:])) * self.dtype.itemsize, self.strides[0] * self.dtype.itemsize
        raise ValueError("unsupported DeviceArray layout for a copy")

    def copy_from(self, src, stream: int | None = None) -> "DeviceArray":
        """Copy from a numpy array (host) or another DeviceArray of the same shape."""
        st = current_stream() if stream is None else stream
        L = _lib.lib()
        set_device(self.device)
        if isinstance(src, DeviceArray):
            assert src.shape == self.shape and src.dtype.itemsize == self.dtype.itemsize
            if self.is_contiguous() and src.is_contiguous():
                _lib.check(L.cotix_memcpy(st, self.data_ptr(), src.data_ptr(), self.nbytes))
            else:
                assert len(self.shape) == 2
                for r in range(self.shape[0]):
                    _lib.check(L.cotix_memcpy(st, self[r].data_ptr(), src[r].data_ptr(),
                                              self.shape[1] * self.dtype.itemsize))
            return self
        a = np.ascontiguousarray(np.asarray(src), dtype=self.dtype)
        assert a.shape == self.shape, (a.shape, self.shape)
        rows, rb, pitch = self._rows()
        for r in range(rows):
            _lib.check(L.cotix_memcpy(st, self.data_ptr() + r * pitch, a.ctypes.data + r * rb, rb))
        _lib.check(L.cotix_stream_synchronize(st))  # `a` may be a temporary
        return self

    def copy_(self, src) -> "DeviceArray":
        return self.copy_from(src)

    def numpy(self) -> np.ndarray:
        out = np.empty(self.shape, self.dtype)
        if self.numel() == 0:
            return out
        st = current_stream()
        L = _lib.lib()
        set_device(self.device)
        rows, rb, pitch = self._rows()
        for r in range(rows):
            _lib.check(L.cotix_memcpy(st, out.ctypes.data + r * rb, self.data_ptr() + r * pitch, rb))
        _lib.check(L.cotix_stream_synchronize(st))
        return out

    def cpu(self):
        return self

    def clone(self) -> "DeviceArray":
        out = DeviceArray.empty(self.shape, self.dtype, self.device,
                                self.strides[0] if len(self.shape) == 2 else None)
        return out.copy_from(self)

    def contiguous(self) -> "DeviceArray":
        if self.is_contiguous():
            return self
        return DeviceArray(self.shape, self.dtype, self.device).copy_from(self)

    def zero_(self) -> "DeviceArray":
        set_device(self.device)
        rows, rb, pitch = self._rows()
        for r in range(rows):
            _lib.check(_lib.lib().cotix_memset(current_stream(), self.data_ptr() + r * pitch, 0, rb))
        return self

    def __repr__(self):
        return f"DeviceArray(shape={self.shape}, dtype={self.dtype}, device={self.device}, strides={self.strides})"
