"""equinox.internal stand-in: eager loops."""
import numpy as _np

from jax import lax as _lax


def while_loop(cond_fun, body_fun, init_val, *, max_steps=None, kind=None, **kw):
    val, steps = init_val, 0
    while bool(_np.asarray(cond_fun(val))) and (max_steps is None or steps < max_steps):
        val = body_fun(val)
        steps += 1
    return val


def scan(f, init, xs, length=None, *, kind=None, **kw):
    return _lax.scan(f, init, xs, length=length)
