"""The reference's own acceptance test for a contact function, restated on batches of shape
records: `_test_contact_info` of test/test_collisions.py:75-159.

For a contact function f(a, b) -> ContactInfo the reference demands, per pair,
    isnan(contact_point)  OR  (contained AND resolved AND minimal AND touching)
  contained : the contact point lies in both shapes                          (:80-82)
  resolved  : after a.move(penetration_vector) the penetration is < eps      (:96-102)
  minimal   : (heavy) no move of length |pen| - 10 eps in any of 20 directions separates,
              unless |pen| < 1.5 eps                                         (:105-121)
  touching  : (heavy) after resolving, some 10-eps nudge re-penetrates by > 5 eps (:124-146)
(`same_edging`, :84-94, calls `_contact_from_edges` with two arguments although it takes six —
that branch cannot run upstream and is not restated.)

`narrow` is any callable (a_recs, b_recs) -> (exists, pen, cp): the CPU oracle in the CPU
tests, the CUDA narrowphase (through the C-ABI) in the GPU tests.  `contains` and `move` are
shape-record helpers (they use the oracle's `contains` restatement as the checker)."""
from __future__ import annotations

import numpy as np


def move(recs: np.ndarray, delta: np.ndarray) -> np.ndarray:
    """shape.move(delta) on records: Circle.position, AABB.lower/upper, Polygon vertices
    (cotix/_convex_shapes.py:31-35,108-111,177-179); polygons are re-sorted like the constructor."""
    from oracle import oracle as o

    r = np.array(recs, np.float32, copy=True)
    d = np.asarray(delta, np.float32).reshape(-1, 2)
    t = r[:, 0].astype(int)
    c = t == 0
    r[c, 4:6] += d[c]
    a = t == 1
    r[a, 4:6] += d[a]
    r[a, 6:8] += d[a]
    p = t == 2
    if p.any():
        n = r[p, 1].astype(int)
        v = r[p, 4:].reshape(-1, 6, 2) + d[p][:, None, :]
        v = v * (np.arange(6)[None, :, None] < n[:, None, None])
        r[p, 4:] = v.reshape(-1, 12)
        r[p] = o.order_clockwise(r[p])
    return r


def check(narrow, a: np.ndarray, b: np.ndarray, heavy: bool = True, small_eps: float = 1e-5):
    """Returns (ok[n], parts) where parts is a dict of the four conditions (bool arrays)."""
    from oracle import oracle as o

    big_eps = 10 * small_eps
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    n = a.shape[0]
    ex, pen, cp = narrow(a, b)[:3]
    nan_cp = np.isnan(cp).any(axis=1)
    cp_q = np.where(nan_cp[:, None], 0.0, cp).astype(np.float32)
    contained = o.contains(a, cp_q) & o.contains(b, cp_q)
    an = move(a, pen)
    _, pen2, _ = narrow(an, b)[:3]
    resolved = np.linalg.norm(pen2, axis=1) < small_eps
    minimal = np.ones(n, bool)
    touching = np.ones(n, bool)
    if heavy:
        ang = np.linspace(0, 2 * np.pi, 20).astype(np.float32)
        dirs = np.stack([np.cos(ang), np.sin(ang)], 1).astype(np.float32)
        length = np.clip(np.linalg.norm(pen, axis=1) - big_eps, 0.0, None).astype(np.float32)
        all_pen = np.ones(n, bool)
        any_deep = np.zeros(n, bool)
        for k in range(20):
            _, pk, _ = narrow(move(a, dirs[k][None, :] * length[:, None]), b)[:3]
            all_pen &= np.linalg.norm(pk, axis=1) > small_eps
            _, qk, _ = narrow(move(an, np.broadcast_to(dirs[k] * np.float32(big_eps), (n, 2))), b)[:3]
            any_deep |= np.linalg.norm(qk, axis=1) > big_eps * 0.5
        minimal = all_pen | (np.linalg.norm(pen, axis=1) < 1.5 * small_eps)
        touching = any_deep
    ok = nan_cp | (contained & resolved & minimal & touching)
    return ok, {"nan": nan_cp, "contained": contained, "resolved": resolved, "minimal": minimal,
                "touching": touching}
