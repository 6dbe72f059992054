"""Pins the CPU oracle (oracle/cotix_oracle.cpp) on everything the reference's own tests hold for
the hot path: the explicit input cases, the exact support expectations, the
input distributions of the randomized tests and — as the acceptance criterion — the reference's
own invariant checker `_test_contact_info` (test/test_collisions.py:75-159, restated in
tests/contact_invariants.py).  CPU only."""
import numpy as np
import pytest

import contact_invariants as ci
import gen_inputs as g
from ref_cases import (AABB_VS_AABB_CASES, AABB_VS_POLYGON_CASES, CIRCLE_VS_AABB_CASES,
                       CIRCLE_VS_POLYGON_CASES, POLY_QUAD, POLY_SQUARE, SUPPORT_CASES, SUPPORT_CIRCLES)


def _all_ok(oracle, a, b, **kw):
    ok, parts = ci.check(oracle.narrowphase, a, b, **kw)
    assert ok.all(), {k: float(v.mean()) for k, v in parts.items()}
    return parts


# ---- explicit cases -------------------------------------------------------------------------
def test_circle_vs_aabb_parametrized(oracle):
    """test/test_collisions.py:226-278 (runs upstream): all 10 known-hard configurations."""
    a = np.stack([oracle.circle_rec(c[0], c[1]) for c in CIRCLE_VS_AABB_CASES])
    b = np.stack([oracle.aabb_rec(c[2], c[3]) for c in CIRCLE_VS_AABB_CASES])
    parts = _all_ok(oracle, a, b)
    assert not parts["nan"].any()  # every one of these configurations is a contact


def test_aabb_vs_aabb_parametrized(oracle):
    """test/test_collisions.py:165-180 (skipped upstream; inputs + invariants)."""
    a = np.stack([oracle.aabb_rec(c[0], c[1]) for c in AABB_VS_AABB_CASES])
    b = np.stack([oracle.aabb_rec(c[2], c[3]) for c in AABB_VS_AABB_CASES])
    _all_ok(oracle, a, b)


def test_aabb_vs_polygon_parametrized(oracle):
    """test/test_collisions.py:378-396 (skipped upstream)."""
    a = np.stack([oracle.aabb_rec(c[0], c[1]) for c in AABB_VS_POLYGON_CASES])
    b = np.stack([oracle.polygon_rec(c[2]) for c in AABB_VS_POLYGON_CASES])
    _all_ok(oracle, a, b, small_eps=1e-3)


def test_circle_vs_polygon_parametrized(oracle):
    """test/test_collisions.py:303-375 (commented out upstream).  The contact point of
    circle_vs_polygon is the nearest point ON a polygon edge (cotix/_contacts.py:166-197) and
    Polygon.contains treats a point on an edge as outside (cotix/_convex_shapes.py:168-175), so
    `contained` cannot hold — which is presumably why the test is disabled upstream.  The three
    geometric conditions (resolved, minimal, touching) do hold."""
    a = np.stack([oracle.circle_rec(c[0], c[1]) for c in CIRCLE_VS_POLYGON_CASES])
    b = np.stack([oracle.polygon_rec(c[2]) for c in CIRCLE_VS_POLYGON_CASES])
    ok, parts = ci.check(oracle.narrowphase, a, b, small_eps=1e-3)
    live = ~parts["nan"]
    assert live.sum() >= 6
    assert (parts["resolved"] & parts["minimal"] & parts["touching"])[live].all()


def test_support_expectations(oracle):
    """test/test_shapes.py:19-35: exact expected supports of a two-circle UniversalShape
    (global support = the part support with the largest projection,
    cotix/_universal_shape.py:45-59)."""
    recs = np.stack([oracle.circle_rec(r, p) for r, p in SUPPORT_CIRCLES])
    for d, want in SUPPORT_CASES:
        dirs = np.repeat(np.float32([d]), len(recs), 0)
        s = oracle.support(recs, dirs)
        best = s[np.argmax(s @ np.float32(d))]
        assert np.allclose(best, want, atol=1e-6), (d, best, want)


def test_support_equivalence_circle(oracle):
    """test/test_shapes.py:8-16: a single-part shape's global support equals the part's."""
    rec = oracle.circle_rec(1.0, [1.0, 1.0])[None]
    for d in [(1, 0), (0, 1), (-1, 0), (0.6, -0.8)]:
        s = oracle.support(rec, np.float32([d]))[0]
        dn = np.float32(d) / np.linalg.norm(d)
        assert np.allclose(s, np.float32([1, 1]) + dn, atol=1e-6)


def test_order_clockwise_consistency(oracle):
    """test/test_geometry_utils.py:7-21: permutation invariance (on <= 6 vertices here)."""
    rng = np.random.default_rng(0)
    for _ in range(200):
        n = int(rng.integers(3, 7))
        v = rng.normal(size=(n, 2)).astype(np.float32)
        base = oracle.polygon_rec(v)
        for _ in range(5):
            again = oracle.polygon_rec(v[rng.permutation(n)])
            assert np.array_equal(base, again)
        # ascending atan2 about the centroid
        vv = base[4:4 + 2 * n].reshape(n, 2)
        c = vv.mean(0)
        ang = np.arctan2(vv[:, 1] - c[1], vv[:, 0] - c[0])
        assert (np.diff(ang) >= -1e-6).all()


def test_simple_world_euler(oracle):
    """test/test_physics_solvers.py:9-40: a ball with velocity (1, 0) drifts right for 100
    steps of dt = 0.1; only the pose changes."""
    px, py, ang = np.float32([0.1]), np.float32([0.1]), np.float32([0.0])
    for _ in range(100):
        px, py, ang = oracle.euler_step(px, py, [1.0], [0.0], ang, [0.0], 0.1)
    assert px[0] > 0.2 and abs(py[0] - 0.1) < 1e-2 and ang[0] == 0.0
    assert abs(px[0] - 10.1) < 1e-3


# ---- randomized property tests with the reference's input distributions ------------------------
@pytest.mark.parametrize("kind,n,heavy", [("cc", 200_000, False), ("cc", 10_000, True),
                                          ("ca", 100_000, False), ("ca", 10_000, True),
                                          ("aa", 100_000, False), ("aa", 10_000, True)])
def test_analytic_contacts_rand(oracle, kind, n, heavy):
    """test_circle_vs_circle_rand :208-223, test_circle_vs_aabb_rand :281-300 (both run upstream),
    test_aabb_vs_aabb_rand :183-205 (skipped upstream): pos ~ N(0,1)^2, r/extent ~ U(0.01, 5)."""
    rng = np.random.default_rng({"cc": 1, "ca": 0, "aa": 2}[kind])
    a = g.random_circles(rng, n) if kind[0] == "c" else g.random_aabbs(rng, n)
    b = g.random_circles(rng, n) if kind[1] == "c" else g.random_aabbs(rng, n)
    parts = _all_ok(oracle, a, b, heavy=heavy)
    assert 0.02 < parts["nan"].mean() < 0.9


@pytest.mark.parametrize("poly", [POLY_SQUARE, POLY_QUAD])
def test_aabb_vs_polygon_rand(oracle, poly):
    """test_aabb_vs_polygon_rand{,_2} :399-433 (skipped upstream), small_eps = 1e-3 as there."""
    rng = np.random.default_rng(3)
    n = 4000
    a = g.random_aabbs(rng, n)
    b = np.repeat(oracle.polygon_rec(poly)[None], n, 0)
    _all_ok(oracle, a, b, small_eps=1e-3)


def test_polygon_vs_polygon_rand(oracle):
    """test_polygon_vs_polygon_rand :436-448 (skipped upstream): random triangle (vertices
    ~ N(0,1)) against the fixed quad; plus convex 3..6-gons of the sweep distribution."""
    rng = np.random.default_rng(4)
    n = 4000
    tri = np.zeros((n, 16), np.float32)
    tri[:, 0], tri[:, 1] = 2, 3
    tri[:, 4:10] = rng.normal(size=(n, 6))
    tri = oracle.order_clockwise(tri)
    b = np.repeat(oracle.polygon_rec(POLY_QUAD)[None], n, 0)
    _all_ok(oracle, tri, b, small_eps=1e-3)
    _all_ok(oracle, g.random_polygons(rng, n), g.random_polygons(rng, n), small_eps=1e-3)


def test_contact_sign_convention(oracle):
    """SURVEY A.2 probe: unit squares overlapping by 0.5 in x -> pen = (-0.5, 0) for A on the
    left: moving A by the penetration vector separates the shapes."""
    a = oracle.polygon_rec([[0, 0], [1, 0], [1, 1], [0, 1]])[None]
    b = oracle.polygon_rec([[0.5, 0], [1.5, 0], [1.5, 1], [0.5, 1]])[None]
    ex, pen, cp, _ = oracle.narrowphase(a, b)
    assert ex[0] == 1 and np.allclose(pen[0], [-0.5, 0.0], atol=1e-6)
    assert 0.5 <= cp[0, 0] <= 1.0 and 0.0 <= cp[0, 1] <= 1.0


def test_miss_is_nan_and_zero_pen(oracle):
    """ContactInfo.nan(): zero penetration vector + NaN contact point (cotix/_contacts.py:19-24)."""
    a = oracle.polygon_rec([[0, 0], [1, 0], [1, 1], [0, 1]])[None]
    b = oracle.polygon_rec([[5, 5], [6, 5], [6, 6]])[None]
    ex, pen, cp, _ = oracle.narrowphase(a, b)
    assert ex[0] == 0 and (pen == 0).all() and np.isnan(cp).all()


def test_gjk_agrees_with_exact_overlap_test(oracle):
    """GJK's yes/no (cotix/_collisions.py:277-310) against an independent float64 separating-axis
    test on convex polygons; disagreements only within 1e-5 of touching."""
    rng = np.random.default_rng(5)
    n = 20000
    a, b = g.random_polygons(rng, n), g.random_polygons(rng, n)
    ex, _ = oracle.gjk(a, b)

    def sat_gap(p, q):
        gap = -np.inf
        for poly in (p, q):
            m = len(poly)
            for k in range(m):
                e = poly[(k + 1) % m] - poly[k]
                ax = np.array([-e[1], e[0]])
                nrm = np.linalg.norm(ax)
                if nrm == 0:
                    continue
                ax = ax / nrm
                pa, pb = p @ ax, q @ ax
                gap = max(gap, pa.min() - pb.max(), pb.min() - pa.max())
        return gap

    bad = 0
    for k in range(0, n, 10):
        p = a[k, 4:4 + 2 * int(a[k, 1])].reshape(-1, 2).astype(np.float64)
        q = b[k, 4:4 + 2 * int(b[k, 1])].reshape(-1, 2).astype(np.float64)
        gap = sat_gap(p, q)
        if abs(gap) < 1e-5:
            continue
        if bool(ex[k]) != (gap < 0):
            bad += 1
    assert bad == 0
