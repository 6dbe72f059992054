"""jax.lax stand-in: eager control flow."""
import numpy as _np

from . import numpy as jnp
from . import tree_util as jtu


def _truth(p):
    return bool(_np.asarray(p))


def cond(pred, true_fn, false_fn, *operands):
    return true_fn(*operands) if _truth(pred) else false_fn(*operands)


def select(pred, a, b):
    return jnp._wrap(_np.where(_np.asarray(pred), _np.asarray(a), _np.asarray(b)))


def _is_arr(x):
    return isinstance(x, _np.ndarray)


def _index(tree, i):
    return jtu.tree_map(lambda x: x[i] if _is_arr(x) else x, tree)


def _length(tree):
    for leaf in jtu.tree_leaves(tree):
        if _is_arr(leaf) and leaf.ndim > 0:
            return leaf.shape[0]
    raise ValueError("no array leaf to map over")


def _stack(outs):
    if not outs:
        return None
    return jtu.tree_map(lambda *xs: jnp.stack(xs) if xs[0] is not None else None, *outs,
                        is_leaf=lambda x: x is None)


def map(f, xs):
    return _stack([f(_index(xs, i)) for i in range(_length(xs))])


def scan(f, init, xs, length=None):
    n = length if xs is None else _length(xs)
    carry, ys = init, []
    for i in range(n):
        carry, y = f(carry, None if xs is None else _index(xs, i))
        ys.append(y)
    return carry, _stack(ys)


def while_loop(cond_fun, body_fun, init_val):
    val = init_val
    while _truth(cond_fun(val)):
        val = body_fun(val)
    return val
