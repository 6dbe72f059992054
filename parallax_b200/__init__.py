"""parallax_b200 — B200-native batched rigid-body step engine for cotix (the package in
DelftMercurians/Parallax).  Hot path only: narrowphase sweep + LunarLander / generic scene step,
hand-written CUDA for sm_100a behind the C-ABI of include/cotix_b200.h.

Importing this package (and every submodule) needs numpy and ctypes only.  Device memory comes
from torch tensors when the caller uses torch, or from `devmem.DeviceArray` (ctypes over the
library's own cudaMalloc wrappers) in a torch-free process such as a JAX program; `jax_ffi`
registers the XLA FFI handlers when jax is importable."""
from . import _lib  # noqa: F401
from ._lib import CotixError, LunarParams, NarrowParams, build, launch_count  # noqa: F401

__all__ = ["_lib", "CotixError", "LunarParams", "NarrowParams", "build", "launch_count"]

_SUBMODULES = ("shapes", "lunar_lander", "sharding", "scene", "viz", "envs", "devmem", "jax_ffi", "_xp")


def __getattr__(name):
    # submodules are imported lazily so that `import parallax_b200` stays cheap
    if name in _SUBMODULES:
        import importlib

        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
