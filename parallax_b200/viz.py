"""Headless stand-in for the reference's painter (cotix/_viz.py).

The reference draws into a pygame window from inside jit through `jax.debug.callback`
(`Painter` :53-75 wrapping `PyPainter` :6-49).  There is no graphics stack on the box, so here the
800 x 600 "screen" is an RGB8 tensor in device memory painted by the CUDA rasteriser
(`cotix_raster_draw`, csrc/raster_kernel.cu); `next()` returns the finished frame instead of
flipping a window.  Same method names, argument meaning and world -> pixel map as the reference.
Pixel-for-pixel agreement with SDL's own line / circle / polygon rasteriser is unpinned (pygame is
not installed anywhere here); the pixel rules are documented in csrc/raster_kernel.cu.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib, _xp
from ._lib import LunarParams

IMG_W, IMG_H, PRIM_WORDS = 800, 600, 16
PRIM_LINE, PRIM_CIRCLE, PRIM_POLYGON = 0, 1, 2
LUNAR_PRIMS = 44


def _f32(v) -> np.float32:
    return np.float32(v)


def scale_vector(vector) -> tuple[int, int]:
    """PyPainter._scale_vector (cotix/_viz.py:33-38): [-8, 8] x [-6, 6] -> [0, 800] x [0, 600], y
    flipped, int() truncation.  float32 inputs are scaled in float32 (what numpy does with the
    arrays a jax callback receives), Python numbers in float64 (the literal pad markers)."""
    x, y = vector[0], vector[1]
    if isinstance(x, (np.floating, np.ndarray)) or isinstance(y, (np.floating, np.ndarray)):
        x, y = _f32(x), _f32(y)
        return (int((x + _f32(8)) / _f32(16) * _f32(800)), int((-y + _f32(6)) / _f32(12) * _f32(600)))
    return (int((x + 8) / 16 * 800), int((-y + 6) / 12 * 600))


def scale_size(size) -> int:
    """PyPainter._scale_size (cotix/_viz.py:40-41)."""
    return int(size * 40)


def _rgb(color) -> int:
    r, g, b = (int(c) & 255 for c in color)
    return (r << 16) | (g << 8) | b


class Painter:
    """`cotix._viz.Painter` without a window: draw_* queue primitives, `next()` rasterises them on
    the GPU onto a black frame and returns it ([600, 800, 3] uint8 device tensor)."""

    def __init__(self, device="cuda"):
        self.device = device
        self._prims: list[list[int]] = []

    def setup(self) -> None:  # PyPainter.setup: nothing to open
        self._prims.clear()

    def draw_circle(self, center, radius, color) -> None:
        cx, cy = scale_vector(center)
        self._prims.append([PRIM_CIRCLE, _rgb(color), scale_size(radius), 0, cx, cy] + [0] * 10)

    def draw_line(self, start, end, color) -> None:
        x0, y0 = scale_vector(start)
        x1, y1 = scale_vector(end)
        self._prims.append([PRIM_LINE, _rgb(color), 0, 0, x0, y0, x1, y1] + [0] * 8)

    def draw_polygon(self, vertices, color) -> None:
        pts = [scale_vector(v) for v in vertices]
        if not 3 <= len(pts) <= 6:
            raise ValueError("draw_polygon takes a convex polygon of 3..6 vertices")
        flat = [c for p in pts for c in p]
        self._prims.append([PRIM_POLYGON, _rgb(color), len(pts), 0] + flat + [0] * (12 - len(flat)))

    def primitives(self) -> np.ndarray:
        """the queued draw calls as records, int32 [n, 16] (include/cotix_b200.h)"""
        return np.asarray(self._prims, dtype=np.int32).reshape(-1, PRIM_WORDS)

    def next(self):
        """PyPainter.next (cotix/_viz.py:43-49): present the frame, start the next one black."""
        prims = self.primitives()
        self._prims = []
        return raster_draw(prims[None], device=self.device)[0]

    def draw_env(self, env) -> None:
        env.draw(self)


def _to_host(x):
    if _xp.is_torch(x):
        return x.detach().cpu().numpy()
    if hasattr(x, "is_native"):
        return x.numpy()
    return np.asarray(x)


def _on_device(x, dtype, xp):
    """numpy / torch / DeviceArray -> contiguous array of `dtype` on backend xp's device"""
    if (xp.name == "torch" and _xp.is_torch(x)) or (xp.name == "native" and hasattr(x, "is_native")):
        return xp.asarray(x, dtype).contiguous()
    return xp.asarray(np.ascontiguousarray(_to_host(x), dtype=dtype), dtype)


def raster_draw(prims, counts=None, device="cuda"):
    """prims int32 [n_img, max_prims, 16] (numpy or device array) -> uint8 [n_img, 600, 800, 3] on
    the device: every image painted in list order onto black (`cotix_raster_draw`)."""
    if _xp.is_torch(prims) and prims.is_cuda or hasattr(prims, "is_native"):
        xp = _xp.backend_of(prims)
    else:
        xp = _xp.backend_for(device)
    prims = _on_device(prims, np.int32, xp)
    n, mp = int(prims.shape[0]), int(prims.shape[1])
    assert prims.dim() == 3 and prims.shape[2] == PRIM_WORDS
    if counts is not None:
        counts = _on_device(counts, np.int32, xp)
    img = xp.empty((n, IMG_H, IMG_W, 3), np.uint8)
    with xp.device_ctx():
        _lib.check(_lib.lib().cotix_raster_draw(
            xp.stream_ptr(), n, prims.data_ptr() if mp else None,
            counts.data_ptr() if counts is not None else None, mp, img.data_ptr()))
        if xp.name == "native":
            xp.synchronize()  # temporaries are dropped on return
    return img


def _index(env_index, xp):
    if env_index is None:
        return None
    return _on_device(env_index, np.int64, xp)


def render_lunar(state, env_index=None, params: LunarParams | None = None):
    """`env.draw(painter); painter.next()` for LunarLander envs (cotix/_lunar_lander.py:220-225):
    uint8 [n_img, 600, 800, 3]; image i shows env env_index[i] (default: every env of `state`)."""
    from . import lunar_lander as ll

    params = params or ll.default_params()
    d = state.data
    xp = _xp.backend_of(d)
    idx = _index(env_index, xp)
    n = int(idx.numel()) if idx is not None else int(d.shape[1])
    img = xp.empty((n, IMG_H, IMG_W, 3), np.uint8)
    with xp.device_ctx():
        _lib.check(_lib.lib().cotix_lunar_draw(
            xp.stream_ptr(), n, d.data_ptr(), d.stride(0),
            idx.data_ptr() if idx is not None else None, img.data_ptr(), C.byref(params)))
        if xp.name == "native":
            xp.synchronize()
    return img


def lunar_primitives(state, env_index=None, params: LunarParams | None = None):
    """the 44 draw calls of LunarLander.draw per env as primitive records, int32 [n_img, 44, 16]"""
    from . import lunar_lander as ll

    params = params or ll.default_params()
    d = state.data
    xp = _xp.backend_of(d)
    idx = _index(env_index, xp)
    n = int(idx.numel()) if idx is not None else int(d.shape[1])
    prims = xp.empty((n, LUNAR_PRIMS, PRIM_WORDS), np.int32)
    with xp.device_ctx():
        _lib.check(_lib.lib().cotix_lunar_prims(
            xp.stream_ptr(), n, d.data_ptr(), d.stride(0),
            idx.data_ptr() if idx is not None else None, prims.data_ptr(), C.byref(params)))
        if xp.name == "native":
            xp.synchronize()
    return prims


def write_ppm(path: str, image) -> None:
    """one frame ([600, 800, 3] uint8, tensor or array) as a binary PPM file"""
    a = _to_host(image)
    assert a.dtype == np.uint8 and a.ndim == 3 and a.shape[2] == 3
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (a.shape[1], a.shape[0]))
        f.write(np.ascontiguousarray(a).tobytes())


def scene_primitives(spec, state, env_index=None) -> np.ndarray:
    """The draw calls of `env.draw(painter)` for a generic scene (`RoboCupEnv.draw`,
    cotix/_robocup.py:140-150) as primitive records, int32 [n_img, n_prims, 16]: per body
    `shape.draw(painter, color)` then `shape.drawEdges(painter, edge_color)`.  An AABB part draws its
    four edges (`AABB.draw` -> `drawEdges`, cotix/_convex_shapes.py:119-133; translated by the body
    position only, :113-117), a circle part a filled circle (:43-44).  `state` is the scene state
    [nb * 6, B] (tensor or array), `spec.draw_list` says what each body paints."""
    s = np.asarray(_to_host(state), np.float32)
    idx = np.arange(s.shape[1]) if env_index is None else np.asarray(env_index, np.int64)
    f = np.float32

    def sx(x):
        return np.trunc(np.nan_to_num((x + f(8)) / f(16) * f(800), nan=-1e6, posinf=1e6, neginf=-1e6).clip(-1048576, 1048576)).astype(np.int32)

    def sy(y):
        return np.trunc(np.nan_to_num((-y + f(6)) / f(12) * f(600), nan=-1e6, posinf=1e6, neginf=-1e6).clip(-1048576, 1048576)).astype(np.int32)

    cols = []

    def line(ax, ay, bx, by, colour):
        q = np.zeros((idx.size, PRIM_WORDS), np.int32)
        q[:, 0], q[:, 1] = PRIM_LINE, _rgb(colour)
        q[:, 4], q[:, 5], q[:, 6], q[:, 7] = sx(ax), sy(ay), sx(bx), sy(by)
        cols.append(q)

    for body, rec, fill, edge in spec.draw_list:
        px, py = s[6 * body][idx].astype(f), s[6 * body + 1][idx].astype(f)
        kind = int(rec[0])
        if kind == 0:  # circle: draw_circle(position, radius, color)
            q = np.zeros((idx.size, PRIM_WORDS), np.int32)
            q[:, 0], q[:, 1], q[:, 2] = PRIM_CIRCLE, _rgb(fill), scale_size(f(rec[2]))
            q[:, 4], q[:, 5] = sx(f(rec[4]) + px), sy(f(rec[5]) + py)
            cols.append(q)
            continue
        lo_x, lo_y, hi_x, hi_y = (f(rec[4]) + px, f(rec[5]) + py, f(rec[6]) + px, f(rec[7]) + py)
        vx, vy = [hi_x, hi_x, lo_x, lo_x], [hi_y, lo_y, lo_y, hi_y]  # AABB.get_edges, :82-93
        for colour in (fill, edge):
            if colour is None:
                continue
            for k in range(4):
                line(vx[k], vy[k], vx[(k + 1) & 3], vy[(k + 1) & 3], colour)
    return np.stack(cols, 1)


def render_scene(spec, state, env_index=None, device="cuda"):
    """frames of a generic scene (e.g. RoboCup), uint8 [n_img, 600, 800, 3] on the device"""
    return raster_draw(scene_primitives(spec, state, env_index), device=device)
