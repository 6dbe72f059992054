"""parallax_b200 — B200-native batched rigid-body step engine for cotix (the package in
DelftMercurians/Parallax).  Hot path only: narrowphase sweep + fused LunarLander step,
hand-written CUDA for sm_100a behind the C-ABI of include/cotix_b200.h."""
from . import _lib  # noqa: F401
from ._lib import CotixError, LunarParams, NarrowParams, build, launch_count  # noqa: F401

__all__ = ["_lib", "CotixError", "LunarParams", "NarrowParams", "build", "launch_count"]


def __getattr__(name):
    # torch-dependent modules are imported lazily so that `import parallax_b200` stays cheap
    if name in ("shapes", "lunar_lander", "sharding", "scene", "viz"):
        import importlib

        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(name)
