"""Synthetic inputs of the BASELINE configs, host-generated with numpy so they do
not depend on any jax version.  Used by tests/, bench.py and __graft_entry__.smoke()."""
from __future__ import annotations

import numpy as np

REC = 16
NF = 36


def random_polygons(rng, n, nverts=None):
    """Convex polygons: n~U{3..6} vertices on a circle r~U(0.5,1.5) about c~N(0,1)^2, already
    in order_clockwise order (ascending atan2 about the centroid, cotix/_geometry_utils.py:60-67)."""
    from oracle import oracle as o

    nv = rng.integers(3, 7, size=n) if nverts is None else np.full(n, nverts)
    c = rng.normal(size=(n, 2)).astype(np.float32)
    r = rng.uniform(0.5, 1.5, size=n).astype(np.float32)
    phi = np.sort(rng.uniform(0, 2 * np.pi, size=(n, 6)), axis=1)
    recs = np.zeros((n, REC), np.float32)
    recs[:, 0] = 2
    recs[:, 1] = nv
    vx = c[:, :1] + r[:, None] * np.cos(phi)
    vy = c[:, 1:] + r[:, None] * np.sin(phi)
    v = np.stack([vx, vy], -1).astype(np.float32)
    mask = (np.arange(6)[None, :] < nv[:, None])
    v = v * mask[..., None]
    recs[:, 4:] = v.reshape(n, 12)
    return o.order_clockwise(recs)


def random_circles(rng, n):
    recs = np.zeros((n, REC), np.float32)
    recs[:, 2] = rng.uniform(0.01, 5.0, size=n)
    recs[:, 4:6] = rng.normal(size=(n, 2))
    return recs


def random_aabbs(rng, n):
    recs = np.zeros((n, REC), np.float32)
    recs[:, 0] = 1
    lo = rng.normal(size=(n, 2)).astype(np.float32)
    recs[:, 4:6] = lo
    recs[:, 6:8] = lo + rng.uniform(0.01, 5.0, size=(n, 2)).astype(np.float32)
    return recs


def random_pairs(rng, n, mix=(0.5, 0.25, 0.125, 0.0625, 0.03125, 0.03125)):
    """BASELINE config 2: 50 % polygon-polygon, 25 % circle-polygon, 12.5 % circle-circle,
    12.5 % circle/AABB-AABB (split 6.25 % circle-AABB, 3.125 % AABB-AABB, 3.125 % AABB-polygon),
    randomly interleaved.  Returns (a[n,16], b[n,16])."""
    kind = rng.choice(6, size=n, p=np.asarray(mix) / np.sum(mix))
    a = np.zeros((n, REC), np.float32)
    b = np.zeros((n, REC), np.float32)
    gens = {
        0: (random_polygons, random_polygons),
        1: (random_circles, random_polygons),
        2: (random_circles, random_circles),
        3: (random_circles, random_aabbs),
        4: (random_aabbs, random_aabbs),
        5: (random_aabbs, random_polygons),
    }
    for k, (ga, gb) in gens.items():
        idx = np.nonzero(kind == k)[0]
        if idx.size:
            a[idx] = ga(rng, idx.size)
            b[idx] = gb(rng, idx.size)
    return a, b


def lunar_keys(B, offset=0):
    """keys (0, i): jax.random.PRNGKey(i) words."""
    return np.stack([np.zeros(B, np.uint32), np.arange(offset, offset + B, dtype=np.uint32)], 1)


def perturbed_lunar_state(state, rng, params):
    """BASELINE config 3: lander pose x~U(-6,6), y~U(-1.5,8), angle~U(-.5,.5), v~N(0,1)^2,
    w~N(0,.5); legs placed rigidly on the lander.  `state` = reset state [36, B] (numpy)."""
    s = state.copy()
    B = s.shape[1]
    x = rng.uniform(-6, 6, B).astype(np.float32)
    y = rng.uniform(-1.5, 8, B).astype(np.float32)
    ang = rng.uniform(-0.5, 0.5, B).astype(np.float32)
    v = rng.normal(size=(2, B)).astype(np.float32)
    w = rng.normal(scale=0.5, size=B).astype(np.float32)
    da = ang - np.float32(params.lander_start[2])
    c, sn = np.cos(da).astype(np.float32), np.sin(da).astype(np.float32)
    s[0], s[1], s[2], s[3], s[4], s[5] = x, y, v[0], v[1], ang, w
    for l in range(2):
        ox, oy = np.float32(params.leg_offset[l][0]), np.float32(params.leg_offset[l][1])
        rx, ry = c * ox - sn * oy, sn * ox + c * oy
        b = 6 * (1 + l)
        s[b + 0], s[b + 1] = x + rx, y + ry
        s[b + 2], s[b + 3] = v[0] - ry * w, v[1] + rx * w
        s[b + 4], s[b + 5] = da, w
    return s
