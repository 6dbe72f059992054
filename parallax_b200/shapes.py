"""Batched host-side mirror of the reference's convex shapes and contact functions.

Same names and argument meaning as the reference (`cotix/_convex_shapes.py` Circle :10-47,
AABB :50-133, Polygon :197-229; `cotix/_contacts.py` ContactInfo :11-27 and the six
`*_vs_*` functions :30-315), but every object is a BATCH of N shapes stored as [N, 16] float32
shape records on the GPU (layout: include/cotix_b200.h), and every contact function is one
launch of the narrowphase kernel through the C-ABI.  Device memory is a torch tensor or a
`devmem.DeviceArray` (`device="native"`, no torch needed — `_xp`); records are assembled on the host
from numpy inputs, on the device from torch inputs.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib, _xp

REC = 16
CIRCLE, AABB_T, POLYGON = 0, 1, 2


def _np_f32(x) -> np.ndarray:
    if _xp.is_torch(x):
        x = x.detach().cpu().numpy()
    elif hasattr(x, "is_native"):
        x = x.numpy()
    return np.asarray(x, dtype=np.float32)


def _device_of(x, device):
    """Inputs that already live on a device keep it; otherwise the `device=` argument decides."""
    if _xp.is_torch(x) and x.is_cuda:
        return x.device
    if hasattr(x, "is_native"):
        return int(x.device)
    return device


class ShapeBatch:
    """N shape records [N, 16] on the device."""

    type_id = -1

    def __init__(self, recs):
        assert recs.shape[-1] == REC and recs.dim() == 2
        self.recs = recs.contiguous()
        self.xp = _xp.backend_of(self.recs)

    def __len__(self):
        return self.recs.shape[0]

    @property
    def device(self):
        return self.recs.device

    def numpy(self) -> np.ndarray:
        return self.xp.to_numpy(self.recs)

    def get_support(self, direction):
        """get_support for one direction per shape -> [N, 2] (reference: get_support)."""
        xp = self.xp
        d = xp.asarray(_np_f32(direction).reshape(-1, 2) if not _same_backend(direction, xp) else direction,
                       np.float32).reshape(-1, 2).contiguous()
        out = xp.empty((d.shape[0], 2), np.float32)
        with xp.device_ctx():
            _lib.check(_lib.lib().cotix_support(xp.stream_ptr(), len(self), self.recs.data_ptr(),
                                                d.data_ptr(), out.data_ptr()))
        return out

    def _rebuilt(self, recs_host: np.ndarray):
        out = type(self).__new__(type(self))
        ShapeBatch.__init__(out, self.xp.asarray(recs_host, np.float32))
        return out


def _same_backend(x, xp) -> bool:
    return (xp.name == "torch" and _xp.is_torch(x)) or (xp.name == "native" and hasattr(x, "is_native"))


def _upload(recs_host: np.ndarray, device):
    return _xp.backend_for(device).asarray(np.ascontiguousarray(recs_host, np.float32), np.float32)


class Circle(ShapeBatch):
    """Circle(radius[N], position[N, 2]) — cotix/_convex_shapes.py:10-47"""

    type_id = CIRCLE

    def __init__(self, radius, position, device="cuda"):
        device = _device_of(radius, device)
        radius = _np_f32(radius).reshape(-1)
        position = _np_f32(position).reshape(-1, 2)
        r = np.zeros((radius.shape[0], REC), np.float32)
        r[:, 0] = CIRCLE
        r[:, 2] = radius
        r[:, 4:6] = position
        super().__init__(_upload(r, device))

    @property
    def radius(self):
        return self.recs[:, 2] if self.xp.name == "torch" else self.numpy()[:, 2]

    @property
    def position(self):
        return self.recs[:, 4:6] if self.xp.name == "torch" else self.numpy()[:, 4:6]

    def move(self, delta):
        r = self.numpy().copy()
        r[:, 4:6] += _np_f32(delta).reshape(-1, 2)
        return self._rebuilt(r)


class AABB(ShapeBatch):
    """AABB(lower[N, 2], upper[N, 2]) — cotix/_convex_shapes.py:50-133"""

    type_id = AABB_T

    def __init__(self, lower, upper, device="cuda"):
        device = _device_of(lower, device)
        lower = _np_f32(lower).reshape(-1, 2)
        upper = _np_f32(upper).reshape(-1, 2)
        r = np.zeros((lower.shape[0], REC), np.float32)
        r[:, 0] = AABB_T
        r[:, 4:6] = lower
        r[:, 6:8] = upper
        super().__init__(_upload(r, device))

    @property
    def lower(self):
        return self.recs[:, 4:6] if self.xp.name == "torch" else self.numpy()[:, 4:6]

    @property
    def upper(self):
        return self.recs[:, 6:8] if self.xp.name == "torch" else self.numpy()[:, 6:8]

    def move(self, delta):
        r = self.numpy().copy()
        d = _np_f32(delta).reshape(-1, 2)
        r[:, 4:6] += d
        r[:, 6:8] += d
        return self._rebuilt(r)


class Polygon(ShapeBatch):
    """Polygon(vertices[N, n, 2]), 3 <= n <= 6 — cotix/_convex_shapes.py:136-229.

    Like the reference constructor, the vertices are re-ordered by `order_clockwise`
    (cotix/_geometry_utils.py:60-67) — on the device.
    """

    type_id = POLYGON

    def __init__(self, vertices, device="cuda", n=None):
        device = _device_of(vertices, device)
        v = _np_f32(vertices)
        if v.ndim == 2:
            v = v[None]
        N, nv = v.shape[0], v.shape[1]
        if not 1 <= nv <= 6:
            raise ValueError("Polygon supports up to 6 vertices")
        r = np.zeros((N, REC), np.float32)
        r[:, 0] = POLYGON
        r[:, 1] = float(nv) if n is None else _np_f32(n)
        r[:, 4:4 + 2 * nv] = v.reshape(N, -1)
        recs = _upload(r, device)
        xp = _xp.backend_of(recs)
        with xp.device_ctx():
            _lib.check(_lib.lib().cotix_order_clockwise(xp.stream_ptr(), N, recs.data_ptr()))
        super().__init__(recs)

    @property
    def vertices(self):
        if not len(self):
            return self.recs[:, 4:4].reshape(0, 0, 2) if self.xp.name == "torch" else np.zeros((0, 0, 2), np.float32)
        if self.xp.name == "torch":
            n = int(self.recs[0, 1].item())
            return self.recs[:, 4:4 + 2 * n].reshape(len(self), n, 2)
        h = self.numpy()
        n = int(h[0, 1])
        return h[:, 4:4 + 2 * n].reshape(len(self), n, 2)

    def move(self, delta):
        d = _np_f32(delta).reshape(-1, 1, 2)
        return Polygon(_np_f32(self.vertices) + d, device=_device_of(self.recs, "cuda"))


def from_records(recs, device="cuda") -> ShapeBatch:
    """Wrap raw [N, 16] records (polygons must already be in order_clockwise order): a numpy array
    is uploaded to `device`; a device array (torch / DeviceArray) is used as it is."""
    if _xp.is_torch(recs) and recs.is_cuda or hasattr(recs, "is_native"):
        return ShapeBatch(recs.reshape(-1, REC))
    return ShapeBatch(_upload(_np_f32(recs).reshape(-1, REC), device))


class ContactInfo:
    """ContactInfo(penetration_vector[N, 2], contact_point[N, 2]) — cotix/_contacts.py:11-27.

    A NaN contact_point means "no contact"; `exists` is its complement as a bool tensor.
    """

    def __init__(self, penetration_vector, contact_point, exists=None, status=None):
        self.penetration_vector = penetration_vector
        self.contact_point = contact_point
        self.exists = exists if exists is not None else ~self._isnan(contact_point)
        self.status = status

    @staticmethod
    def _isnan(cp):
        if _xp.is_torch(cp):
            import torch

            return torch.isnan(cp).any(-1)
        return np.isnan(cp.numpy() if hasattr(cp, "is_native") else np.asarray(cp)).any(-1)

    def isnan(self):
        return self._isnan(self.contact_point)

    def invert(self):
        pv = self.penetration_vector
        if hasattr(pv, "is_native"):
            pv = _xp.backend_of(pv).asarray(-pv.numpy(), np.float32)
        else:
            pv = -pv
        return ContactInfo(pv, self.contact_point, self.exists, self.status)

    def numpy(self):
        """(exists, penetration_vector, contact_point, status) as host numpy arrays."""
        def h(x):
            if x is None:
                return None
            return x.numpy() if hasattr(x, "is_native") else (x.detach().cpu().numpy() if _xp.is_torch(x) else np.asarray(x))
        return h(self.exists), h(self.penetration_vector), h(self.contact_point), h(self.status)


SWEEP_PIPELINE_MIN_PAIRS = 65536
_sweep_ws: dict = {}


def _sweep_workspace(xp, n: int):
    key = (xp.name, xp.device_index(), n)
    ws = _sweep_ws.get(key)
    if ws is None:
        nbytes = int(_lib.lib().cotix_narrowphase_workspace_bytes(n))
        if any(k[0] == "native" for k in _sweep_ws):
            xp.synchronize()
        _sweep_ws.clear()
        with xp.device_ctx():
            ws = xp.empty((nbytes + 256,), np.uint8)
        _sweep_ws[key] = ws
    return ws


def narrowphase(a: ShapeBatch, b: ShapeBatch, params: _lib.NarrowParams | None = None,
                mode: str = "auto") -> ContactInfo:
    """One contact query per pair (a[k], b[k]) — the sweep (BASELINE config 2).
    mode: "single" = one mixed-type kernel (cotix_narrowphase), "pipeline" = classify / GJK / EPA
    passes over type-homogeneous queues (cotix_narrowphase_ws), "auto" = pipeline from 65536 pairs.
    Identical results."""
    if len(a) != len(b):
        raise ValueError("shape batches differ in length")
    xp = a.xp
    n = len(a)
    params = params or _lib.default_narrow_params()
    exists = xp.empty((n,), np.uint8)
    pen = xp.empty((n, 2), np.float32)
    cp = xp.empty((n, 2), np.float32)
    status = xp.empty((n,), np.uint8)
    use_pipe = mode == "pipeline" or (mode == "auto" and n >= SWEEP_PIPELINE_MIN_PAIRS)
    with xp.device_ctx():
        args = (xp.stream_ptr(), n, a.recs.data_ptr(), b.recs.data_ptr(), exists.data_ptr(), pen.data_ptr(),
                cp.data_ptr(), status.data_ptr(), C.byref(params))
        if use_pipe and n > 0:
            ws = _sweep_workspace(xp, n)
            wp = (ws.data_ptr() + 255) & ~255
            _lib.check(_lib.lib().cotix_narrowphase_ws(*args, wp, ws.numel() - (wp - ws.data_ptr())))
        else:
            _lib.check(_lib.lib().cotix_narrowphase(*args))
    return ContactInfo(pen, cp, exists.bool() if xp.name == "torch" else exists, status)


def _typed(fn_name, ta, tb):
    def f(a: ShapeBatch, b: ShapeBatch, params=None) -> ContactInfo:
        if a.type_id != ta or b.type_id != tb:
            raise TypeError(f"{fn_name} expects ({ta}, {tb}) shape types")
        return narrowphase(a, b, params)

    f.__name__ = fn_name
    f.__doc__ = f"Batched `{fn_name}` of cotix/_contacts.py (one pair per row)."
    return f


circle_vs_circle = _typed("circle_vs_circle", CIRCLE, CIRCLE)
aabb_vs_aabb = _typed("aabb_vs_aabb", AABB_T, AABB_T)
circle_vs_aabb = _typed("circle_vs_aabb", CIRCLE, AABB_T)
circle_vs_polygon = _typed("circle_vs_polygon", CIRCLE, POLYGON)
aabb_vs_polygon = _typed("aabb_vs_polygon", AABB_T, POLYGON)
polygon_vs_polygon = _typed("polygon_vs_polygon", POLYGON, POLYGON)


def narrowphase_host(a_recs: np.ndarray, b_recs: np.ndarray, params=None, device: int = 0):
    """HOST-buffer entry point (cotix_narrowphase_host): numpy in, numpy out."""
    a = np.ascontiguousarray(a_recs, np.float32).reshape(-1, REC)
    b = np.ascontiguousarray(b_recs, np.float32).reshape(-1, REC)
    n = a.shape[0]
    params = params or _lib.default_narrow_params()
    exists = np.empty(n, np.uint8)
    pen = np.empty((n, 2), np.float32)
    cp = np.empty((n, 2), np.float32)
    status = np.empty(n, np.uint8)
    _lib.check(_lib.lib().cotix_narrowphase_host(device, n, a.ctypes.data, b.ctypes.data, exists.ctypes.data,
                                                 pen.ctypes.data, cp.ctypes.data, status.ctypes.data,
                                                 C.byref(params)))
    return exists, pen, cp, status
