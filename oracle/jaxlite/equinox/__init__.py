"""equinox stand-in (see ../README.md): Modules as plain pytrees, tree_at, eager filter_* helpers."""
import functools

import numpy as _np

import jax.tree_util as jtu
from jax import lax as _lax

from . import internal  # noqa: F401


class _FieldSpec:
    def __init__(self, static=False, converter=None, default=None, has_default=False):
        self.static, self.converter, self.default, self.has_default = static, converter, default, has_default


_MISSING = object()


def field(*, static=False, converter=None, default=_MISSING, **kw):
    return _FieldSpec(static, converter, None if default is _MISSING else default, default is not _MISSING)


class AbstractVar:
    def __class_getitem__(cls, item):
        return cls


class AbstractClassVar(AbstractVar):
    pass


class Module:
    __jaxlite_module__ = True

    def __init_subclass__(cls, strict=False, **kw):
        super().__init_subclass__(**kw)
        fields, specs = [], {}
        for klass in reversed(cls.__mro__):
            for name, ann in getattr(klass, "__annotations__", {}).items():
                if name.startswith("__"):
                    continue
                if name not in fields:
                    fields.append(name)
                v = klass.__dict__.get(name, _MISSING)
                if isinstance(v, _FieldSpec):
                    specs[name] = v
                elif v is not _MISSING and not callable(v) and not isinstance(v, property):
                    specs[name] = _FieldSpec(default=v, has_default=True)
        cls.__jaxlite_fields__ = fields
        cls.__jaxlite_specs__ = specs
        user_init = any("__init__" in k.__dict__ for k in cls.__mro__ if k not in (Module, object))
        if not user_init:
            def __init__(self, *args, **kwargs):
                names = type(self).__jaxlite_fields__
                sp = type(self).__jaxlite_specs__
                vals = dict(zip(names, args))
                vals.update(kwargs)
                for n in names:
                    if n in vals:
                        object.__setattr__(self, n, vals[n])
                    elif n in sp and sp[n].has_default:
                        object.__setattr__(self, n, sp[n].default)
                    else:
                        raise TypeError(f"{type(self).__name__}: missing field {n}")
            cls.__init__ = __init__
        orig = cls.__init__

        @functools.wraps(orig)
        def wrapped(self, *a, **k):
            orig(self, *a, **k)
            if type(self) is cls:  # outermost constructor: apply converters once
                for n, sp in type(self).__jaxlite_specs__.items():
                    if sp.converter is not None and n in self.__dict__:
                        object.__setattr__(self, n, sp.converter(self.__dict__[n]))

        cls.__init__ = wrapped

    def __jaxlite_dynamic_fields__(self):
        sp = type(self).__jaxlite_specs__
        return [n for n in type(self).__jaxlite_fields__ if n in self.__dict__ and not (n in sp and sp[n].static)]

    def __hash__(self):
        return id(self)

    def __eq__(self, other):
        return self is other

    def __repr__(self):
        return f"{type(self).__name__}({', '.join(f'{k}=...' for k in self.__dict__)})"


def is_array(x):
    return isinstance(x, _np.ndarray)


def is_inexact_array(x):
    return isinstance(x, _np.ndarray) and x.dtype.kind == "f"


def if_array(axis):
    return lambda x: axis if is_array(x) else None


def filter_jit(f=None, **kw):
    if f is None:
        return lambda g: g
    return f


def filter_vmap(fn=None, *, in_axes=None, out_axes=None, **kw):
    if fn is None:
        return lambda g: filter_vmap(g, in_axes=in_axes, out_axes=out_axes)

    def g(*args):
        n = _lax._length(args)
        return _lax._stack([fn(*_lax._index(args, i)) for i in range(n)])

    return g


# eqx.error_if: "raise" (equinox's default, EQX_ON_ERROR=raise) or "record" — the value passes
# through and the message is appended to ERROR_LOG (what the engine's status bits stand for)
ERROR_MODE = "raise"
ERROR_LOG = []


def error_if(x, pred, msg=""):
    if bool(_np.any(_np.asarray(pred))):
        if ERROR_MODE == "raise":
            raise RuntimeError(msg)
        ERROR_LOG.append(msg)
    return x


class _Sentinel:
    __slots__ = ("i",)

    def __init__(self, i):
        self.i = i


def tree_at(where, pytree, replace=_MISSING, replace_fn=None, is_leaf=None):
    """Out-of-place replacement of the node(s) `where` selects (a single node or a tuple/list of nodes)."""
    leaves, treedef = jtu.tree_flatten(pytree)
    shadow = jtu.tree_unflatten(treedef, [_Sentinel(i) for i in range(len(leaves))])
    sel = where(shadow)
    multi = isinstance(sel, (tuple, list)) and not jtu._is_module(sel) and all(
        isinstance(s, _Sentinel) or jtu._children(s) is not None for s in sel) and _selected_is_sequence(sel, shadow)
    targets = list(sel) if multi else [sel]
    repl = list(replace) if multi and replace is not _MISSING else [replace]

    def rec(sh, orig):
        for t, r in zip(targets, repl if replace is not _MISSING else [None] * len(targets)):
            if sh is t:
                return replace_fn(orig) if replace is _MISSING else r
        c = jtu._children(sh)
        if c is None:
            return orig
        ch, rebuild_shadow = c
        och, rebuild = jtu._children(orig)
        return rebuild([rec(a, b) for a, b in zip(ch, och)])

    return rec(shadow, pytree)


def _selected_is_sequence(sel, shadow):
    """True when `sel` is a fresh tuple/list of nodes, not itself a node of the tree."""
    found = []

    def rec(x):
        if x is sel:
            found.append(True)
        c = jtu._children(x)
        if c is not None:
            for v in c[0]:
                rec(v)

    rec(shadow)
    return not found


def partition(pytree, filter_spec, replace=None):
    f = filter_spec if callable(filter_spec) else (lambda x: bool(filter_spec))
    a = jtu.tree_map(lambda x: x if f(x) else replace, pytree)
    b = jtu.tree_map(lambda x: replace if f(x) else x, pytree)
    return a, b


def combine(*trees):
    def pick(*xs):
        for x in xs:
            if x is not None:
                return x
        return None

    return jtu.tree_map(pick, *trees, is_leaf=lambda x: x is None)


def apply_updates(model, updates):
    return jtu.tree_map(lambda m, u: m if u is None else m + u, model, updates, is_leaf=lambda x: x is None)
