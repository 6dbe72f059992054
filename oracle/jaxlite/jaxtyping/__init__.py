"""jaxtyping stand-in: annotations only (no runtime checking)."""


class _Ann:
    def __class_getitem__(cls, item):
        return cls


class Array:
    pass


class Float(_Ann):
    pass


class Int(_Ann):
    pass


class Bool(_Ann):
    pass


class Shaped(_Ann):
    pass


class PyTree(_Ann):
    pass


class Key(_Ann):
    pass


PRNGKeyArray = Array
ArrayLike = Array
Scalar = Array


def jaxtyped(f=None, **kw):
    if f is None:
        return lambda g: g
    return f


def install_import_hook(*a, **k):
    class _Ctx:
        def __enter__(self):
            return self

        def __exit__(self, *e):
            return False

    return _Ctx()
