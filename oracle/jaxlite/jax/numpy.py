"""jax.numpy stand-in: eager numpy, float32 / int32 defaults, one IEEE operation at a time,
left-to-right sums, sin/cos/atan2 from the engine's cx_trig.h definitions (through the CPU
checker's `trig` export)."""
import builtins as _bi

import numpy as _np

pi = float(_np.pi)
nan = float("nan")
inf = float("inf")
float32 = _np.float32
int32 = _np.int32
uint32 = _np.uint32
bool_ = _np.bool_
newaxis = None


class _At:
    def __init__(self, arr):
        self.arr = arr

    def __getitem__(self, idx):
        return _AtIdx(self.arr, idx)


class _AtIdx:
    def __init__(self, arr, idx):
        self.arr, self.idx = arr, idx

    def _idx(self):
        idx = self.idx
        if isinstance(idx, tuple):
            return tuple(int(i) if isinstance(i, _np.ndarray) and i.ndim == 0 else i for i in idx)
        if isinstance(idx, _np.ndarray) and idx.ndim == 0:
            return int(idx)
        return idx

    def set(self, value):
        out = _np.array(self.arr, copy=True)
        out[self._idx()] = _np.asarray(value).astype(out.dtype)
        return _wrap(out)

    def add(self, value):
        out = _np.array(self.arr, copy=True)
        out[self._idx()] = (out[self._idx()] + _np.asarray(value)).astype(out.dtype)
        return _wrap(out)


class JArr(_np.ndarray):
    """ndarray with jax's `.at[...]` functional updates; results stay float32 / int32."""

    @property
    def at(self):
        return _At(self)

    def __array_finalize__(self, obj):
        pass

    def __iter__(self):
        for i in range(self.shape[0]):
            yield self[i]

    def __getitem__(self, idx):
        if isinstance(idx, tuple):
            idx = tuple(int(i) if isinstance(i, _np.ndarray) and i.ndim == 0 and i.dtype.kind in "iu" else i
                        for i in idx)
        # jax clamps out-of-bounds retrieval indices (cotix/_colliders.py:217 relies on it)
        if isinstance(idx, tuple) and _bi.all(isinstance(i, (int, _np.integer)) for i in idx) and len(idx) <= self.ndim:
            idx = tuple(_bi.min(_bi.max(int(i), -n), n - 1) for i, n in zip(idx, self.shape))
        elif isinstance(idx, (int, _np.integer)) and self.ndim >= 1:
            idx = _bi.min(_bi.max(int(idx), -self.shape[0]), self.shape[0] - 1)
        r = _np.ndarray.__getitem__(self, idx)
        if not isinstance(r, _np.ndarray):
            r = _np.asarray(r).view(JArr)
        return r

    def __array_ufunc__(self, ufunc, method, *inputs, out=None, **kwargs):
        # jax's weak-type promotion: python floats and integer arrays meeting a float become float32
        has_float = _bi.any(isinstance(v, float) or (isinstance(v, _np.ndarray) and v.dtype.kind == "f") or
                        isinstance(v, _np.floating) for v in inputs)
        conv = []
        for v in inputs:
            if isinstance(v, JArr):
                v = v.view(_np.ndarray)
            if isinstance(v, float):
                v = _np.float32(v)
            elif has_float and isinstance(v, _np.ndarray) and v.dtype.kind in "iub":
                v = v.astype(_np.float32)
            elif has_float and isinstance(v, (_np.integer, _np.bool_)):
                v = _np.float32(v)
            elif isinstance(v, _np.ndarray) and v.dtype == _np.float64:
                v = v.astype(_np.float32)
            conv.append(v)
        with _np.errstate(all="ignore"):
            r = getattr(ufunc, method)(*conv, **kwargs)
        if isinstance(r, tuple):
            return tuple(_wrap(v) for v in r)
        return _wrap(r)

    def astype(self, dtype, **k):
        return _wrap(_np.ndarray.astype(self.view(_np.ndarray), dtype))

    def __hash__(self):
        return id(self)

    def block_until_ready(self):
        return self


def _wrap(x):
    a = _np.asarray(x)
    if a.dtype == _np.float64:
        a = a.astype(_np.float32)
    elif a.dtype == _np.int64:
        a = a.astype(_np.int32)
    return a.view(JArr)


def array(x, dtype=None):
    if isinstance(x, (list, tuple)):
        x = [_np.asarray(v) for v in x]
    a = _np.array(x)
    if dtype is not None:
        a = a.astype(dtype)
    return _wrap(a)


asarray = array


def zeros(shape, dtype=float32):
    return _wrap(_np.zeros(shape, dtype))


def ones(shape, dtype=float32):
    return _wrap(_np.ones(shape, dtype))


def zeros_like(x):
    return _wrap(_np.zeros_like(_np.asarray(x)))


def nans(shape):
    return _wrap(_np.full(shape, _np.nan, _np.float32))


def arange(*a, **k):
    return _wrap(_np.arange(*a, **k))


def linspace(*a, **k):
    return _wrap(_np.linspace(*a, **k))


def _seq_sum(a, axis):
    """sum along `axis`, strictly left to right"""
    a = _np.moveaxis(_np.asarray(a), axis, 0)
    acc = a[0].copy() if a.shape[0] else _np.zeros(a.shape[1:], a.dtype)
    for k in range(1, a.shape[0]):
        acc = acc + a[k]
    return acc


def sum(x, axis=None):
    a = _np.asarray(x)
    if a.dtype == _np.bool_:
        a = a.astype(_np.int32)
    if axis is None:
        return _wrap(_seq_sum(a.reshape(-1), 0))
    return _wrap(_seq_sum(a, axis))


def mean(x, axis=None):
    a = _np.asarray(x)
    n = a.size if axis is None else a.shape[axis]
    return _wrap(_np.asarray(sum(a, axis)) / _np.float32(n))


def dot(a, b):
    a, b = _np.asarray(a), _np.asarray(b)
    if a.dtype.kind != "f" and b.dtype.kind == "f":
        a = a.astype(b.dtype)
    if b.dtype.kind != "f" and a.dtype.kind == "f":
        b = b.astype(a.dtype)
    if a.ndim == 1 and b.ndim == 1:
        return _wrap(_seq_sum(a * b, 0))
    if a.ndim == 2 and b.ndim == 1:
        return _wrap(_seq_sum(a * b[None, :], 1))
    if a.ndim == 2 and b.ndim == 2:
        return _wrap(_seq_sum(a[:, :, None] * b[None, :, :], 1))
    if a.ndim == 1 and b.ndim == 2:
        return _wrap(_seq_sum(a[:, None] * b, 0))
    raise NotImplementedError("dot shapes")


def cross(a, b):
    a, b = _np.asarray(a), _np.asarray(b)
    assert a.shape[-1] == 2 and b.shape[-1] == 2
    return _wrap(a[..., 0] * b[..., 1] - a[..., 1] * b[..., 0])


class _Linalg:
    @staticmethod
    def norm(x, axis=None):
        a = _np.asarray(x)
        return _wrap(_np.sqrt(_np.asarray(sum(a * a, axis))))

    @staticmethod
    def inv(m):
        return _wrap(_np.linalg.inv(_np.asarray(m)).astype(_np.float32))


linalg = _Linalg()


def _trig(x, y=None):
    from oracle import oracle as _o

    xa = _np.ascontiguousarray(_np.asarray(x, _np.float32).reshape(-1))
    ya = xa if y is None else _np.ascontiguousarray(_np.asarray(y, _np.float32).reshape(-1))
    return _o.trig(xa, ya)


def sin(x):
    return _wrap(_trig(x)[0].reshape(_np.shape(x)))


def cos(x):
    return _wrap(_trig(x)[1].reshape(_np.shape(x)))


def arctan2(y, x):
    return _wrap(_trig(x, y)[2].reshape(_np.shape(x)))


def clip(x, a_min=None, a_max=None, min=None, max=None):
    lo = a_min if a_min is not None else min
    hi = a_max if a_max is not None else max
    r = _np.asarray(x)
    if lo is not None:
        r = _np.maximum(r, _np.asarray(lo, r.dtype) if r.dtype.kind == "f" else lo)
    if hi is not None:
        r = _np.minimum(r, _np.asarray(hi, r.dtype) if r.dtype.kind == "f" else hi)
    return _wrap(r)


def argsort(x, axis=-1):
    return _wrap(_np.argsort(_np.asarray(x), axis=axis, kind="stable"))


def stack(xs, axis=0):
    return _wrap(_np.stack([_np.asarray(v) for v in xs], axis=axis))


def concatenate(xs, axis=0):
    return _wrap(_np.concatenate([_np.asarray(v) for v in xs], axis=axis))


def _fwd(name):
    f = getattr(_np, name)

    def g(*a, **k):
        a = [(_np.asarray(v) if isinstance(v, (_np.ndarray, list, tuple)) else v) for v in a]
        with _np.errstate(all="ignore"):
            return _wrap(f(*a, **k))

    g.__name__ = name
    return g


for _n in ("all", "any", "isnan", "maximum", "minimum", "logical_or", "logical_not", "logical_and",
           "logical_xor", "argmin", "argmax", "sign", "roll", "where", "equal", "sqrt", "abs", "square",
           "arctan", "arccos", "isinf", "isfinite", "transpose", "reshape", "max", "min", "floor", "exp"):
    globals()[_n] = _fwd(_n)


def nonzero(x, size=None, fill_value=None):
    r = _np.nonzero(_np.asarray(x))
    return tuple(_wrap(v) for v in r)
