"""Batched host-side mirror of the reference's convex shapes and contact functions.

Same names and argument meaning as the reference (`cotix/_convex_shapes.py` Circle :10-47,
AABB :50-133, Polygon :197-229; `cotix/_contacts.py` ContactInfo :11-27 and the six
`*_vs_*` functions :30-315), but every object is a BATCH of N shapes stored as [N, 16] float32
shape records on the GPU (layout: include/cotix_b200.h), and every contact function is one
launch of the narrowphase kernel through the C-ABI.  torch is used only to own device memory
and streams.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib

REC = 16
CIRCLE, AABB_T, POLYGON = 0, 1, 2


def _dev_f32(x, device):
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=torch.float32)
    return torch.as_tensor(np.asarray(x, dtype=np.float32), device=device)


def _stream_ptr(device) -> int:
    return torch.cuda.current_stream(device).cuda_stream


class ShapeBatch:
    """N shape records [N, 16] on the device."""

    type_id = -1

    def __init__(self, recs: torch.Tensor):
        assert recs.is_cuda and recs.dtype == torch.float32 and recs.shape[-1] == REC
        self.recs = recs.contiguous()

    def __len__(self):
        return self.recs.shape[0]

    @property
    def device(self):
        return self.recs.device

    def get_support(self, direction):
        """get_support for one direction per shape -> [N, 2] (reference: get_support)."""
        d = _dev_f32(direction, self.device).reshape(-1, 2).contiguous()
        out = torch.empty_like(d)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().cotix_support(_stream_ptr(self.device), len(self), self.recs.data_ptr(),
                                                d.data_ptr(), out.data_ptr()))
        return out


class Circle(ShapeBatch):
    """Circle(radius[N], position[N, 2]) — cotix/_convex_shapes.py:10-47"""

    type_id = CIRCLE

    def __init__(self, radius, position, device="cuda"):
        radius = _dev_f32(radius, device).reshape(-1)
        position = _dev_f32(position, device).reshape(-1, 2)
        r = torch.zeros((radius.shape[0], REC), dtype=torch.float32, device=radius.device)
        r[:, 0] = CIRCLE
        r[:, 2] = radius
        r[:, 4:6] = position
        super().__init__(r)

    @property
    def radius(self):
        return self.recs[:, 2]

    @property
    def position(self):
        return self.recs[:, 4:6]

    def move(self, delta):
        out = Circle.__new__(Circle)
        r = self.recs.clone()
        r[:, 4:6] += _dev_f32(delta, self.device).reshape(-1, 2)
        ShapeBatch.__init__(out, r)
        return out


class AABB(ShapeBatch):
    """AABB(lower[N, 2], upper[N, 2]) — cotix/_convex_shapes.py:50-133"""

    type_id = AABB_T

    def __init__(self, lower, upper, device="cuda"):
        lower = _dev_f32(lower, device).reshape(-1, 2)
        upper = _dev_f32(upper, device).reshape(-1, 2)
        r = torch.zeros((lower.shape[0], REC), dtype=torch.float32, device=lower.device)
        r[:, 0] = AABB_T
        r[:, 4:6] = lower
        r[:, 6:8] = upper
        super().__init__(r)

    @property
    def lower(self):
        return self.recs[:, 4:6]

    @property
    def upper(self):
        return self.recs[:, 6:8]

    def move(self, delta):
        out = AABB.__new__(AABB)
        r = self.recs.clone()
        d = _dev_f32(delta, self.device).reshape(-1, 2)
        r[:, 4:6] += d
        r[:, 6:8] += d
        ShapeBatch.__init__(out, r)
        return out


class Polygon(ShapeBatch):
    """Polygon(vertices[N, n, 2]), 3 <= n <= 6 — cotix/_convex_shapes.py:136-229.

    Like the reference constructor, the vertices are re-ordered by `order_clockwise`
    (cotix/_geometry_utils.py:60-67) — on the device.
    """

    type_id = POLYGON

    def __init__(self, vertices, device="cuda", n=None):
        v = _dev_f32(vertices, device)
        if v.dim() == 2:
            v = v[None]
        N, nv = v.shape[0], v.shape[1]
        if not 1 <= nv <= 6:
            raise ValueError("Polygon supports up to 6 vertices")
        r = torch.zeros((N, REC), dtype=torch.float32, device=v.device)
        r[:, 0] = POLYGON
        r[:, 1] = float(nv) if n is None else _dev_f32(n, device)
        r[:, 4:4 + 2 * nv] = v.reshape(N, -1)
        with torch.cuda.device(r.device):
            _lib.check(_lib.lib().cotix_order_clockwise(_stream_ptr(r.device), N, r.data_ptr()))
        super().__init__(r)

    @property
    def vertices(self):
        n = int(self.recs[0, 1].item()) if len(self) else 0
        return self.recs[:, 4:4 + 2 * n].reshape(len(self), n, 2)

    def move(self, delta):
        d = _dev_f32(delta, self.device).reshape(-1, 1, 2)
        return Polygon(self.vertices + d, device=self.device)


def from_records(recs, device="cuda") -> ShapeBatch:
    """Wrap raw [N, 16] records (polygons must already be in order_clockwise order)."""
    return ShapeBatch(_dev_f32(recs, device).reshape(-1, REC))


class ContactInfo:
    """ContactInfo(penetration_vector[N, 2], contact_point[N, 2]) — cotix/_contacts.py:11-27.

    A NaN contact_point means "no contact"; `exists` is its complement as a bool tensor.
    """

    def __init__(self, penetration_vector, contact_point, exists=None, status=None):
        self.penetration_vector = penetration_vector
        self.contact_point = contact_point
        self.exists = exists if exists is not None else ~torch.isnan(contact_point).any(-1)
        self.status = status

    def isnan(self):
        return torch.isnan(self.contact_point).any(-1)

    def invert(self):
        return ContactInfo(-self.penetration_vector, self.contact_point, self.exists, self.status)


SWEEP_PIPELINE_MIN_PAIRS = 65536
_sweep_ws: dict = {}


def _sweep_workspace(device, n: int) -> torch.Tensor:
    key = (device.index, n)
    ws = _sweep_ws.get(key)
    if ws is None:
        nbytes = int(_lib.lib().cotix_narrowphase_workspace_bytes(n))
        _sweep_ws.clear()
        ws = torch.empty(nbytes + 256, dtype=torch.uint8, device=device)
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
    dev = a.device
    n = len(a)
    params = params or _lib.default_narrow_params()
    exists = torch.empty(n, dtype=torch.uint8, device=dev)
    pen = torch.empty((n, 2), dtype=torch.float32, device=dev)
    cp = torch.empty((n, 2), dtype=torch.float32, device=dev)
    status = torch.empty(n, dtype=torch.uint8, device=dev)
    use_pipe = mode == "pipeline" or (mode == "auto" and n >= SWEEP_PIPELINE_MIN_PAIRS)
    with torch.cuda.device(dev):
        args = (_stream_ptr(dev), n, a.recs.data_ptr(), b.recs.data_ptr(), exists.data_ptr(), pen.data_ptr(),
                cp.data_ptr(), status.data_ptr(), C.byref(params))
        if use_pipe and n > 0:
            ws = _sweep_workspace(dev, n)
            wp = (ws.data_ptr() + 255) & ~255
            _lib.check(_lib.lib().cotix_narrowphase_ws(*args, wp, ws.numel() - (wp - ws.data_ptr())))
        else:
            _lib.check(_lib.lib().cotix_narrowphase(*args))
    return ContactInfo(pen, cp, exists.bool(), status)


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
