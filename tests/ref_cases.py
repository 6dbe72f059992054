"""Inputs the reference's own tests hold for the hot path (test/test_collisions.py), copied as
DATA (numbers only) so that the CPU oracle and the CUDA path can be checked on them without jax.
"""
# test/test_collisions.py:229-272 — test_circle_vs_aabb_parametrized (this test RUNS upstream):
# (radius, circle position, aabb lower, aabb upper)
CIRCLE_VS_AABB_CASES = [
    (4.808976, (0.52343243, 0.38244677), (1.2948408, 1.4734308), (3.3397233, 6.3817973)),
    (1.0, (1.0, 1.0), (-2.0, -2.0), (0.4, 0.7)),
    (5.0, (0.0, 0.0), (-2.0, -2.0), (2.0, 2.0)),
    (3.7427633, (-0.0277214, 1.0449156), (-0.6238362, -1.1297362), (1.3405488, -0.5544366)),
    (0.5361439, (-0.4457733, 0.5882554), (-0.44587463, -0.73396504), (0.0717122, 3.0028129)),
    (1.0, (0.0, 0.0), (-2.0, -2.0), (2.0, 2.0)),
    (0.01, (0.0, 1.8), (-2.0, -2.0), (2.0, 2.0)),
    (1.0, (0.1, 0.2), (-2.0, -2.3), (2.0, 2.0)),
    (1.0, (-0.3, 0.05), (-2.1, -2.3), (2.0, 2.0)),
    (1.0, (-0.12, -0.56), (-2.0, -2.0), (2.2, 2.3)),
]
# test/test_collisions.py:168-177 — test_aabb_vs_aabb_parametrized (skipped upstream; inputs only)
AABB_VS_AABB_CASES = [
    ((-0.72877413, 0.4557484), (2.8047075, 1.2977818), (0.54837185, -0.6289791), (2.0380392, 2.3360252)),
]
# test/test_collisions.py:381-391 — test_aabb_vs_polygon_parametrized (skipped upstream)
AABB_VS_POLYGON_CASES = [
    ((0.13606448, -2.6069396), (0.39154065, 0.7682829), [[0.5, 0.5], [-0.5, -0.5], [0.5, -0.5], [-0.5, 0.5]]),
    ((-0.8, -0.8), (-0.4, -0.4), [[0.0, 1.0], [-0.5, -0.5], [0.5, -0.5]]),
]
# test/test_collisions.py:307-334 — test_circle_vs_polygon_parametrized (commented out upstream)
_TRI = [[0.0, 1.0], [-0.5, -0.5], [0.5, -0.5]]
CIRCLE_VS_POLYGON_CASES = [
    (2.3251324, (-0.9682081, 0.6601708), _TRI),
    (0.9460892, (-0.27002454, -0.07793757), _TRI),
    (1.8268951, (-1.0627501, -0.18848151), _TRI),
    (1.9288263, (0.96713173, -0.5197397), _TRI),
    (3.82366, (0.89924145, 0.10981558), _TRI),
    (3.3993943, (-0.5438884, -0.28964934), _TRI),
    (1.0, (0.0, 0.0), [[0.0, 0.9], [2.0, 3.0], [-2.0, 3.0]]),
]
# fixed polygons used by the (skipped) randomized polygon tests, :410-457
POLY_SQUARE = [[0.5, 0.5], [-0.5, -0.5], [0.5, -0.5], [-0.5, 0.5]]
POLY_QUAD = [[0.3, 0.556], [-0.1, -0.2], [0.4, -0.3], [-0.8, 1.5]]
# test/test_shapes.py:19-35 — exact support expectations of a two-circle UniversalShape
SUPPORT_CASES = [  # (direction, expected support)
    ((1.0, 0.0), (2.0, 1.0)),
    ((-1.0, 0.0), (-10.5, 0.0)),
    ((0.0, 1.0), (1.0, 2.0)),
    ((0.0, -1.0), (-10.0, -0.5)),
]
SUPPORT_CIRCLES = [(0.5, (-10.0, 0.0)), (1.0, (1.0, 1.0))]
