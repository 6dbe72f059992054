"""jax.random stand-in: the published algorithms of jax's default (non-partitionable) threefry2x32
PRNG — independent numpy implementation (the C++ checker has its own)."""
import numpy as _np

from . import numpy as jnp

_ROT = ((13, 15, 26, 6), (17, 29, 16, 24))


def _rotl(x, r):
    return ((x << _np.uint32(r)) | (x >> _np.uint32(32 - r))).astype(_np.uint32)


def _threefry2x32(k0, k1, x0, x1):
    with _np.errstate(over="ignore"):
        k0, k1 = _np.uint32(k0), _np.uint32(k1)
        ks = (k0, k1, _np.uint32(k0 ^ k1 ^ _np.uint32(0x1BD11BDA)))
        x0 = (_np.asarray(x0, _np.uint32) + ks[0]).astype(_np.uint32)
        x1 = (_np.asarray(x1, _np.uint32) + ks[1]).astype(_np.uint32)
        for g in range(5):
            for r in _ROT[g % 2]:
                x0 = (x0 + x1).astype(_np.uint32)
                x1 = _rotl(x1, r)
                x1 = x1 ^ x0
            x0 = (x0 + ks[(g + 1) % 3]).astype(_np.uint32)
            x1 = (x1 + ks[(g + 2) % 3] + _np.uint32(g + 1)).astype(_np.uint32)
    return x0, x1


def _bits(key, n):
    key = _np.asarray(key).view(_np.ndarray).astype(_np.uint32).reshape(2)
    m = n + (n & 1)
    ctr = _np.arange(m, dtype=_np.uint32)
    ctr[n:] = 0
    h = m // 2
    o0, o1 = _threefry2x32(key[0], key[1], ctr[:h], ctr[h:])
    return _np.concatenate([o0, o1])[:n]


def PRNGKey(seed):
    seed = int(seed)
    return jnp._wrap(_np.array([(seed >> 32) & 0xFFFFFFFF, seed & 0xFFFFFFFF], _np.uint32))


key = PRNGKey


def split(key, num=2):
    if isinstance(num, tuple):
        assert len(num) == 1
        num = num[0]
    num = int(num)
    return jnp._wrap(_bits(key, 2 * num).reshape(num, 2))


def _size(shape):
    if shape is None or shape == ():
        return 1, ()
    if isinstance(shape, int):
        shape = (shape,)
    return int(_np.prod(shape)), tuple(shape)


def uniform(key, shape=(), dtype=_np.float32, minval=0.0, maxval=1.0):
    n, shp = _size(shape)
    bits = _bits(key, n)
    f = ((bits >> _np.uint32(9)) | _np.uint32(0x3F800000)).view(_np.float32) - _np.float32(1.0)
    lo, hi = _np.float32(minval), _np.float32(maxval)
    v = f * (hi - lo) + lo
    return jnp._wrap(_np.maximum(lo, v).astype(_np.float32).reshape(shp))


def _erf_inv(x):
    """XLA's float32 erf_inv (Giles' polynomial), evaluated with separate multiplies and adds."""
    x = _np.asarray(x, _np.float32)
    w = -_np.log1p((-x * x).astype(_np.float32)).astype(_np.float32)
    lt = w < 5.0
    w1 = (w - _np.float32(2.5)).astype(_np.float32)
    w2 = (_np.sqrt(w) - _np.float32(3.0)).astype(_np.float32)
    c1 = [2.81022636e-08, 3.43273939e-07, -3.5233877e-06, -4.39150654e-06, 0.00021858087, -0.00125372503,
          -0.00417768164, 0.246640727, 1.50140941]
    c2 = [-0.000200214257, 0.000100950558, 0.00134934322, -0.00367342844, 0.00573950773, -0.0076224613,
          0.00943887047, 1.00167406, 2.83297682]
    ww = _np.where(lt, w1, w2).astype(_np.float32)
    p = _np.where(lt, _np.float32(c1[0]), _np.float32(c2[0])).astype(_np.float32)
    for a, b in zip(c1[1:], c2[1:]):
        p = (_np.where(lt, _np.float32(a), _np.float32(b)) + (p * ww).astype(_np.float32)).astype(_np.float32)
    return (p * x).astype(_np.float32)


def normal(key, shape=(), dtype=_np.float32):
    n, shp = _size(shape)
    lo = _np.nextafter(_np.float32(-1.0), _np.float32(0.0))
    u = _np.asarray(uniform(key, (n,), minval=lo, maxval=1.0))
    return jnp._wrap((_np.float32(_np.sqrt(2.0)) * _erf_inv(u)).astype(_np.float32).reshape(shp))


def bernoulli(key, p=0.5, shape=None):
    n, shp = _size(shape if shape is not None else ())
    return jnp._wrap(_np.asarray(uniform(key, (n,))) < _np.float32(p)).reshape(shp)


def choice(key, a, shape=(), replace=True, p=None):
    a = _np.asarray(a)
    assert shape == () and p is not None
    p = _np.asarray(p, _np.float32)
    c = _np.cumsum(p, dtype=_np.float32)
    u = _np.asarray(uniform(key, ()))
    r = c[-1] * (_np.float32(1.0) - u)
    ind = int(_np.searchsorted(c, r, side="left"))
    return jnp._wrap(a[min(ind, a.shape[0] - 1)])
