"""Minimal pytrees: tuples, lists, dicts, None and equinox Modules are nodes; everything else
(arrays, python scalars, callables) is a leaf."""
import functools

import numpy as np


class Partial(functools.partial):
    pass


def _is_module(x):
    return hasattr(type(x), "__jaxlite_module__")


def _children(x):
    """(children list, rebuild function) or None for a leaf"""
    if x is None:
        return [], (lambda ch: None)
    if isinstance(x, tuple) and not hasattr(x, "_fields"):
        return list(x), (lambda ch: tuple(ch))
    if isinstance(x, list):
        return list(x), (lambda ch: list(ch))
    if isinstance(x, dict):
        keys = list(x.keys())
        return [x[k] for k in keys], (lambda ch: dict(zip(keys, ch)))
    if _is_module(x):
        names = x.__jaxlite_dynamic_fields__()
        vals = [getattr(x, n) for n in names]

        def rebuild(ch, x=x, names=names):
            new = object.__new__(type(x))
            for k, v in x.__dict__.items():
                object.__setattr__(new, k, v)
            for n, v in zip(names, ch):
                object.__setattr__(new, n, v)
            return new

        return vals, rebuild
    return None


def tree_flatten(tree, is_leaf=None):
    leaves = []

    def rec(x):
        if is_leaf is not None and is_leaf(x):
            leaves.append(x)
            return ("leaf",)
        c = _children(x)
        if c is None:
            leaves.append(x)
            return ("leaf",)
        ch, rebuild = c
        return ("node", rebuild, [rec(v) for v in ch])

    return leaves, rec(tree)


def tree_unflatten(treedef, leaves):
    it = iter(leaves)

    def rec(d):
        if d[0] == "leaf":
            return next(it)
        return d[1]([rec(c) for c in d[2]])

    return rec(treedef)


def tree_leaves(tree, is_leaf=None):
    return tree_flatten(tree, is_leaf)[0]


def tree_map(f, tree, *rest, is_leaf=None):
    def rec(x, others):
        if is_leaf is not None and is_leaf(x):
            return f(x, *others)
        c = _children(x)
        if c is None:
            return f(x, *others)
        ch, rebuild = c
        och = []
        for o in others:
            oc = _children(o)
            if oc is None or len(oc[0]) != len(ch):
                raise ValueError("tree_map: tree structures differ")
            och.append(oc[0])
        return rebuild([rec(v, [o[i] for o in och]) for i, v in enumerate(ch)])

    return rec(tree, list(rest))


def tree_reduce(f, tree, initializer=None, is_leaf=None):
    leaves = tree_leaves(tree, is_leaf)
    if initializer is None:
        return functools.reduce(f, leaves)
    return functools.reduce(f, leaves, initializer)
