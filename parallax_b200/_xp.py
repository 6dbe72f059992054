"""Array backends of the host-side mirror: torch tensors (when torch is importable and the caller
uses them) or `devmem.DeviceArray` (no torch anywhere).  The mirror only needs to allocate, copy
and take pointers — the library's kernels do everything else — so a backend is a dozen functions.
torch is imported lazily and ONLY when a torch object or a torch device string is handed in."""
from __future__ import annotations

import contextlib

import numpy as np

from . import devmem


def is_torch(x) -> bool:
    return type(x).__module__.split(".")[0] == "torch"


def torch_available() -> bool:
    try:
        import torch  # noqa: F401

        return True
    except Exception:
        return False


_T2N = None


def _torch_dtype(dt):
    import torch

    return {np.dtype(np.float32): torch.float32, np.dtype(np.uint8): torch.uint8, np.dtype(np.int32): torch.int32,
            np.dtype(np.uint32): torch.int32, np.dtype(np.int64): torch.int64, np.dtype(np.float64): torch.float64,
            np.dtype(np.bool_): torch.bool}[np.dtype(dt)]


class TorchBackend:
    name = "torch"

    def __init__(self, device):
        import torch

        self.torch = torch
        self.device = torch.device(device)
        if self.device.type == "cuda" and self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())

    def empty(self, shape, dtype=np.float32, row_stride=None):
        t = self.torch
        shape = tuple(shape) if hasattr(shape, "__len__") else (shape,)
        if row_stride is not None and len(shape) >= 2 and row_stride != shape[-1]:
            st = [1, int(row_stride)]
            for s in reversed(shape[1:-1]):
                st.append(st[-1] * int(s))
            return t.empty_strided(shape, tuple(reversed(st)), dtype=_torch_dtype(dtype), device=self.device)
        return t.empty(shape, dtype=_torch_dtype(dtype), device=self.device)

    def zeros(self, shape, dtype=np.float32, row_stride=None):
        return self.empty(shape, dtype, row_stride).zero_()

    def asarray(self, x, dtype=np.float32):
        """numpy / torch (any device) -> tensor of `dtype` on this device (no copy when already so)."""
        t = self.torch
        if is_torch(x):
            return x.to(device=self.device, dtype=_torch_dtype(dtype))
        a = np.ascontiguousarray(np.asarray(x), dtype=np.dtype(dtype))
        if a.dtype == np.uint32:
            a = a.view(np.int32)
        return t.from_numpy(a).to(self.device)

    def to_numpy(self, x) -> np.ndarray:
        return x.detach().cpu().numpy()

    def stream_ptr(self) -> int:
        return self.torch.cuda.current_stream(self.device).cuda_stream

    def device_index(self) -> int:
        return self.device.index

    def device_ctx(self):
        return self.torch.cuda.device(self.device)

    def synchronize(self):
        self.torch.cuda.synchronize(self.device)


class NativeBackend:
    name = "native"

    def __init__(self, device: int = 0):
        self.device = int(device)

    def empty(self, shape, dtype=np.float32, row_stride=None):
        return devmem.DeviceArray.empty(shape, dtype, self.device, row_stride)

    def zeros(self, shape, dtype=np.float32, row_stride=None):
        return devmem.DeviceArray.zeros(shape, dtype, self.device, row_stride)

    def asarray(self, x, dtype=np.float32):
        if isinstance(x, devmem.DeviceArray):
            assert x.dtype.itemsize == np.dtype(dtype).itemsize
            return x
        if is_torch(x):
            x = x.detach().cpu().numpy()
        return devmem.DeviceArray.from_numpy(np.ascontiguousarray(np.asarray(x), dtype=np.dtype(dtype)), self.device)

    def to_numpy(self, x) -> np.ndarray:
        return x.numpy()

    def stream_ptr(self) -> int:
        return devmem.current_stream()

    def device_index(self) -> int:
        return self.device

    @contextlib.contextmanager
    def device_ctx(self):
        devmem.set_device(self.device)
        yield

    def synchronize(self):
        devmem.synchronize()


def backend_of(x):
    """The backend that owns array `x`."""
    if isinstance(x, devmem.DeviceArray):
        return NativeBackend(x.device)
    if is_torch(x):
        return TorchBackend(x.device)
    raise TypeError(f"expected a torch tensor or a DeviceArray, got {type(x).__name__}")


def backend_for(device):
    """Backend for a `device=` argument: "native" / "native:1" / an int -> DeviceArray backend;
    "cuda" / "cuda:1" / a torch.device -> torch.  "cuda" falls back to the native backend when torch
    is not importable (a torch-free process)."""
    if isinstance(device, (int, np.integer)):
        return NativeBackend(int(device))
    if isinstance(device, str):
        if device.startswith("native"):
            return NativeBackend(int(device.split(":")[1]) if ":" in device else 0)
        if not torch_available():
            return NativeBackend(int(device.split(":")[1]) if ":" in device else 0)
        return TorchBackend(device)
    if is_torch(device):  # a torch.device
        return TorchBackend(device)
    raise TypeError(f"bad device {device!r}")


def row_ld(t, what: str) -> int:
    """Leading dimension of a [rows, B] buffer whose rows are contiguous (stride(1) == 1)."""
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{what}: rows must be contiguous (stride(1) == 1); call .contiguous()")
    return int(t.stride(0)) if t.shape[0] > 1 else max(int(t.stride(0)), int(t.shape[1]))


def with_ld(xp, t, rows: int, B: int, ld: int, what: str):
    """`t` ([rows, B], any backend / numpy) as a float32 buffer of THIS backend whose leading
    dimension is `ld` — the C-ABI takes ONE ld for every [.][ld] buffer of a call."""
    if t is None:
        return None
    t = xp.asarray(t, np.float32)
    assert tuple(t.shape) == (rows, B), f"{what}: expected shape {(rows, B)}, got {tuple(t.shape)}"
    if t.stride(1) == 1 and (rows == 1 or t.stride(0) == ld):
        return t
    out = xp.empty((rows, B), np.float32, ld)
    out.copy_(t)
    return out
