"""K6 — the headless rasteriser that stands in for the pygame window of cotix/_viz.py.

pygame is not installed (and there is no graphics stack), so agreement with SDL's rasteriser is
unpinned; what is checked: the world -> pixel map against the reference's formula
(cotix/_viz.py:33-41), the oracle's sequential painter against hand-computed pixel sets, and the
CUDA kernels (cotix_raster_draw / cotix_lunar_draw / cotix_lunar_prims) bit for bit against the
oracle on random primitive lists and on LunarLander states."""
import numpy as np
import pytest

from gen_inputs import lunar_keys, perturbed_lunar_state

W, H = 800, 600


def _line(x0, y0, x1, y1, rgb=0xFFFFFF):
    return [0, rgb, 0, 0, x0, y0, x1, y1] + [0] * 8


def _circle(cx, cy, r, rgb=0xC80000):
    return [1, rgb, r, 0, cx, cy] + [0] * 10


def _poly(pts, rgb=0x00FF00):
    flat = [c for p in pts for c in p]
    return [2, rgb, len(pts), 0] + flat + [0] * (12 - len(flat))


def _lit(img):
    return {(int(x), int(y)) for y, x in zip(*np.nonzero(img.any(-1)))}


# ---- CPU: the oracle's painter and the host-side scale functions -----------------------------------
def test_oracle_painter_known_pixels(oracle):
    o = oracle
    img = o.raster_draw(np.asarray([[_line(10, 5, 14, 5)]], np.int32))[0]
    assert _lit(img) == {(x, 5) for x in range(10, 15)}
    assert tuple(img[5, 10]) == (255, 255, 255)
    # 45 degrees, and a shallow line: minor coordinate rounds half up from the smaller end
    assert _lit(o.raster_draw(np.asarray([[_line(3, 3, 6, 6)]], np.int32))[0]) == {(3, 3), (4, 4), (5, 5), (6, 6)}
    assert _lit(o.raster_draw(np.asarray([[_line(0, 0, 4, 2)]], np.int32))[0]) == {(0, 0), (1, 1), (2, 1), (3, 2), (4, 2)}
    # direction does not matter (the line is drawn from its smaller end)
    a = o.raster_draw(np.asarray([[_line(17, 40, 230, 99)]], np.int32))
    b = o.raster_draw(np.asarray([[_line(230, 99, 17, 40)]], np.int32))
    assert np.array_equal(a, b) and len(_lit(a[0])) == 214
    # steep line: one pixel per row
    assert len(_lit(o.raster_draw(np.asarray([[_line(50, 10, 53, 110)]], np.int32))[0])) == 101
    # circle: dx^2 + dy^2 <= r^2; nothing for r < 1
    assert _lit(o.raster_draw(np.asarray([[_circle(20, 20, 1)]], np.int32))[0]) == {(20, 20), (19, 20), (21, 20), (20, 19), (20, 21)}
    assert len(_lit(o.raster_draw(np.asarray([[_circle(100, 100, 4)]], np.int32))[0])) == 49
    assert not _lit(o.raster_draw(np.asarray([[_circle(20, 20, 0)]], np.int32))[0])
    # filled triangle, boundary included: (0,0) (4,0) (0,4) -> x + y <= 4
    assert _lit(o.raster_draw(np.asarray([[_poly([(0, 0), (4, 0), (0, 4)])]], np.int32))[0]) == {
        (x, y) for x in range(5) for y in range(5) if x + y <= 4}
    # painter's order: the later primitive wins; counts cut the list; clipping drops off-screen pixels
    two = np.asarray([[_line(0, 7, 20, 7, 0xFFFFFF), _line(10, 0, 10, 20, 0xFF0000)]], np.int32)
    assert tuple(o.raster_draw(two)[0][7, 10]) == (255, 0, 0)
    assert tuple(o.raster_draw(two[:, ::-1])[0][7, 10]) == (255, 255, 255)
    assert tuple(o.raster_draw(two, counts=[1])[0][0, 10]) == (0, 0, 0)
    far = o.raster_draw(np.asarray([[_line(-5000, 300, 5000, 300), _circle(-2, -2, 4), _poly([(790, 590), (900, 590), (790, 700)])]], np.int32))[0]
    lit = _lit(far)
    assert {(x, 300) for x in range(W)} <= lit and (0, 0) in lit and (799, 599) in lit


def test_scale_vector_is_the_reference_map():
    from parallax_b200 import viz

    # cotix/_viz.py:33-41 on Python numbers (the pad markers of cotix/_lunar_lander.py:223-224)
    for v in [(-2, -1.8), (-2, -1.0), (2, -1.8), (2, -1.0), (0, 0), (-8, 6), (7.99, -5.99)]:
        assert viz.scale_vector(v) == (int((v[0] + 8) / 16 * 800), int((-v[1] + 6) / 12 * 600))
    assert viz.scale_vector((0, 0)) == (400, 300) and viz.scale_vector((-8, 6)) == (0, 0)
    # float32 arrays (what a jax callback hands to PyPainter) are scaled in float32
    rng = np.random.default_rng(0)
    for v in rng.uniform(-9, 9, size=(200, 2)).astype(np.float32):
        ex = (int((v[0] + 8) / 16 * 800), int((-v[1] + 6) / 12 * 600))  # numpy float32 scalar arithmetic
        assert viz.scale_vector(v) == ex
    assert viz.scale_size(0.1) == 4 and viz.scale_size(np.float32(0.5)) == 20


def test_robocup_draw_calls_follow_the_reference(oracle):
    """RoboCupEnv.draw (cotix/_robocup.py:132-150): per body shape.draw(colour) then drawEdges(edge colour);
    AABBs paint their four edges, the ball a filled circle; robots (new) Circle.draw's grey."""
    from parallax_b200 import scene as sc
    from parallax_b200 import viz

    spec = sc.robocup_scene(n_robots=2, areas_collide=False)
    B = 3
    st = np.repeat(spec.init_state[:, :, None], B, axis=2).astype(np.float32).reshape(spec.nb * 6, B)
    st[6 * 4, 1] = 1.5          # the ball of env 1 moved
    st[6 * 4 + 1, 1] = -2.0
    pr = viz.scene_primitives(spec, st)
    # 2 areas + 3 + 3 goal boxes, each 4 edges in the fill colour and 4 in the edge colour; ball; 2 robots
    assert pr.shape == (B, 8 * 8 + 1 + 2, 16)
    sv = lambda x, y: (int((np.float32(x) + 8) / 16 * 800), int((-np.float32(y) + 6) / 12 * 600))
    # field: upper = (5.2, 3.7), first edge (upper) -> (upper.x, lower.y)   (AABB.get_edges, _convex_shapes.py:82-93)
    assert list(pr[0, 0, :8]) == [0, 0x00B400, 0, 0, *sv(5.2, 3.7), *sv(5.2, -3.7)]
    assert list(pr[0, 4, :2]) == [0, 0xFFFFFF] and list(pr[0, 4, 4:8]) == list(pr[0, 0, 4:8])
    assert list(pr[0, 16, :2]) == [0, 0xFFFF00] and list(pr[0, 40, :2]) == [0, 0x0080FF]
    ball = pr[:, 64]
    assert list(ball[0, :6]) == [1, 0xFF0000, int(np.float32(0.022) * np.float32(3) * 40), 0, 400, 300]
    assert list(ball[1, 4:6]) == list(sv(1.5, -2.0))
    assert list(pr[0, 65, :3]) == [1, 0x808080, int(0.09 * 40)]
    img = oracle.raster_draw(pr)
    assert tuple(img[0, 300, 400]) == (255, 0, 0) and tuple(img[1, 300, 400]) == (0, 0, 0)
    assert tuple(img[0, 115, 400]) == (255, 255, 255)      # the field's white top edge over the green one
    assert tuple(img[0, 300, 175]) == (255, 255, 0) or (img[0, :, 170:185] == (255, 255, 0)).all(-1).any()


# ---- GPU: the CUDA rasteriser against the oracle, bit for bit -----------------------------------------
def _random_prims(rng, n_img, max_prims):
    prims = np.zeros((n_img, max_prims, 16), np.int32)
    colours = [0xFFFFFF, 0xFF0000, 0xC80000, 0x00FF00, 0x0000FF, 0x808080, 0x000000]
    for i in range(n_img):
        for k in range(max_prims):
            rgb = colours[rng.integers(len(colours))]
            kind = rng.integers(4)
            if kind == 0:      # line, mostly on screen
                p = rng.integers(-60, 860, size=4)
                prims[i, k] = _line(int(p[0]), int(p[1] * 0.75), int(p[2]), int(p[3] * 0.75), rgb)
            elif kind == 1:    # line with far-away end points, steep and shallow, degenerate
                c = rng.integers(5)
                if c == 0:
                    p = rng.integers(-6000, 6000, size=4)
                    prims[i, k] = _line(*(int(v) for v in p), rgb)
                elif c == 1:
                    x = int(rng.integers(0, W)); prims[i, k] = _line(x, int(rng.integers(-900, 0)), x + int(rng.integers(-3, 4)), int(rng.integers(600, 1500)), rgb)
                elif c == 2:
                    y = int(rng.integers(0, H)); prims[i, k] = _line(int(rng.integers(-2000, 0)), y, int(rng.integers(800, 3000)), y + int(rng.integers(-3, 4)), rgb)
                elif c == 3:
                    x, y = int(rng.integers(0, W)), int(rng.integers(0, H)); prims[i, k] = _line(x, y, x, y, rgb)
                else:
                    prims[i, k] = _line(-(1 << 20), int(rng.integers(0, H)), 1 << 20, int(rng.integers(0, H)), rgb)
            elif kind == 2:
                prims[i, k] = _circle(int(rng.integers(-20, 820)), int(rng.integers(-20, 620)), int(rng.integers(0, 60)), rgb)
            else:              # convex polygon: points on a circle in angular order, either winding
                nv = int(rng.integers(3, 7))
                ang = np.sort(rng.uniform(0, 2 * np.pi, nv))
                if rng.integers(2):
                    ang = ang[::-1]
                cx, cy, r = rng.integers(-50, 850), rng.integers(-50, 650), rng.integers(1, 150)
                pts = [(int(cx + r * np.cos(a)), int(cy + r * np.sin(a))) for a in ang]
                prims[i, k] = _poly(pts, rgb)
    return prims


@pytest.mark.gpu
def test_raster_draw_matches_the_oracle(cuda_lib, oracle):
    from parallax_b200 import viz

    rng = np.random.default_rng(7)
    prims = _random_prims(rng, 9, 40)
    counts = np.asarray([40, 0, 1, 17, 40, 39, 5, 40, 23], np.int32)
    img = viz.raster_draw(prims, counts).cpu().numpy()
    ref = oracle.raster_draw(prims, counts)
    assert np.array_equal(img, ref)
    assert not img[1].any() and img[0].any()
    # lists longer than one staging chunk (64 records), with a count that ends inside a chunk
    big = _random_prims(rng, 2, 150)
    cb = np.asarray([150, 70], np.int32)
    assert np.array_equal(viz.raster_draw(big, cb).cpu().numpy(), oracle.raster_draw(big, cb))
    # without counts every record is drawn; an empty list gives black frames; zero images is fine
    assert np.array_equal(viz.raster_draw(prims[:3]).cpu().numpy(), oracle.raster_draw(prims[:3]))
    assert not viz.raster_draw(np.zeros((2, 0, 16), np.int32)).cpu().numpy().any()
    assert viz.raster_draw(np.zeros((0, 4, 16), np.int32)).shape == (0, H, W, 3)


def _lunar_prims_numpy(o, s, P):
    """numpy float32 restatement of LunarLander.draw's draw calls (cotix/_lunar_lander.py:220-225,
    cotix/_convex_shapes.py:160-163,189-194, cotix/_viz.py:33-38) for state s [36, B]"""
    f = np.float32
    B = s.shape[1]
    out = np.zeros((B, 44, 16), np.int32)

    def to_int(v):
        v = np.asarray(v, np.float32)
        r = np.where(np.isnan(v), f(-(1 << 20)), np.clip(v, f(-1048576.0), f(1048576.0)))
        return np.trunc(r).astype(np.int64).astype(np.int32)

    def sx(x):
        return to_int((x + f(8)) / f(16) * f(800))

    def sy(y):
        return to_int((-y + f(6)) / f(12) * f(600))

    def put(k, ax, ay, bx, by, rgb=0xFFFFFF):
        out[:, k, 0], out[:, k, 1] = 0, rgb
        out[:, k, 4], out[:, k, 5], out[:, k, 6], out[:, k, 7] = sx(ax), sy(ay), sx(bx), sy(by)

    locs = [np.asarray(P.lander_local, f).reshape(6, 2), np.asarray(P.leg_local[0], f).reshape(4, 2),
            np.asarray(P.leg_local[1], f).reshape(4, 2)]
    k = 0
    for b, loc in enumerate(locs):
        px, py, ang = s[6 * b], s[6 * b + 1], s[6 * b + 4]
        sn, cs, _ = o.trig(ang, ang)
        wx = [cs * v[0] + (-sn) * v[1] + px * f(1) for v in loc]
        wy = [sn * v[0] + cs * v[1] + py * f(1) for v in loc]
        n = len(loc)
        for i in range(n):
            j = n - 1 if i == 0 else i - 1
            put(k, wx[i], wy[i], wx[j], wy[j])
            k += 1
    X = [np.full(B, -P.terrain_x_outer, f), s[26], s[27], np.full(B, -P.pad_half_width, f),
         np.full(B, P.pad_half_width, f), s[28], s[29], np.full(B, P.terrain_x_outer, f)]
    fl = np.full(B, P.ground_floor, f)
    for g in range(7):
        x0, x1, h0, h1 = X[g], X[g + 1], s[18 + g], s[19 + g]
        put(k, x0, h0, x1, h1); put(k + 1, x1, h1, x1, fl); put(k + 2, x1, fl, x0, fl); put(k + 3, x0, fl, x0, h0)
        k += 4
    for x in (-2, 2):  # the literal pad markers, Python float arithmetic
        out[:, k, 0], out[:, k, 1] = 0, 0xFF0000
        out[:, k, 4], out[:, k, 5] = int((x + 8) / 16 * 800), int((1.8 + 6) / 12 * 600)
        out[:, k, 6], out[:, k, 7] = int((x + 8) / 16 * 800), int((1.0 + 6) / 12 * 600)
        k += 1
    assert k == 44
    return out


@pytest.mark.gpu
def test_lunar_draw_matches_the_reference_draw_calls(cuda_lib, oracle, exact_trig):
    import torch

    from parallax_b200 import lunar_lander as ll
    from parallax_b200 import viz

    o = oracle
    P = ll.default_params()
    B = 96
    s = perturbed_lunar_state(o.lunar_reset(lunar_keys(B), P), np.random.default_rng(2), P)
    s[1, :8] += np.float32(30.0)      # some landers far off screen
    s[0, 8] = np.nan                  # and a broken one: its lines are dropped
    st = ll.LunarState.from_numpy(s)
    for _ in range(3):
        st, _, _ = ll.step(st, None, P)
    s = st.numpy()
    prims = viz.lunar_primitives(st, params=P).cpu().numpy()
    assert np.array_equal(prims, _lunar_prims_numpy(o, s, P))
    img = viz.render_lunar(st, params=P).cpu().numpy()
    assert np.array_equal(img, o.raster_draw(prims))
    # reset pose: the lander is on screen, the pad markers are red, the background black
    st0 = ll.reset(lunar_keys(4), P)
    f0 = viz.render_lunar(st0, params=P).cpu().numpy()
    assert tuple(f0[0, 370, 300]) == (255, 0, 0) and tuple(f0[0, 370, 500]) == (255, 0, 0) and tuple(f0[0, 0, 0]) == (0, 0, 0)
    assert tuple(f0[0, 400, 400]) == (255, 255, 255)                      # the pad's top edge, y = -2
    assert (f0[0] == 255).all(-1).sum() > 300
    # env_index picks and repeats envs
    idx = np.asarray([5, 5, 90, 0], np.int64)
    sub = viz.render_lunar(st, idx, P).cpu().numpy()
    assert np.array_equal(sub, img[idx])
    assert viz.render_lunar(st, np.zeros(0, np.int64), P).shape == (0, H, W, 3)


@pytest.mark.gpu
def test_painter_mirror(cuda_lib, oracle, tmp_path):
    """the reference's Painter API (cotix/_viz.py:53-75): draw_* then next()"""
    from parallax_b200 import viz

    p = viz.Painter()
    p.setup()
    p.draw_circle(np.asarray([0.5, 1.0], np.float32), 0.1, (200, 0, 0))   # examples/test_viz.py:33-35
    p.draw_polygon([(-1, -1), (1, -1), (0, 1)], (0, 255, 0))
    p.draw_line((-2, -1.8), (-2, -1.0), color=(255, 0, 0))                # cotix/_lunar_lander.py:223
    prims = p.primitives()
    assert prims.shape == (3, 16) and list(prims[2, :8]) == [0, 0xFF0000, 0, 0, 300, 390, 300, 350]
    assert list(prims[0, :6]) == [1, 0xC80000, 4, 0, 425, 250]
    frame = p.next()
    assert np.array_equal(frame.cpu().numpy(), oracle.raster_draw(prims[None])[0])
    assert p.primitives().shape == (0, 16) and not p.next().any()         # next() starts a black frame
    viz.write_ppm(str(tmp_path / "f.ppm"), frame)
    raw = open(tmp_path / "f.ppm", "rb").read()
    assert raw.startswith(b"P6\n800 600\n255\n") and len(raw) == 15 + W * H * 3


@pytest.mark.gpu
def test_raster_argument_checks(cuda_lib):
    import torch

    from parallax_b200 import _lib

    img = torch.empty(W * H * 3 + 16, dtype=torch.uint8, device="cuda")
    pr = torch.zeros((1, 2, 16), dtype=torch.int32, device="cuda")
    L = _lib.lib()
    with pytest.raises(_lib.CotixError):
        _lib.check(L.cotix_raster_draw(None, 1, pr.data_ptr(), None, 2, None))
    with pytest.raises(_lib.CotixError):
        _lib.check(L.cotix_raster_draw(None, 1, pr.data_ptr(), None, 2, img.data_ptr() + 1))
    with pytest.raises(_lib.CotixError):
        _lib.check(L.cotix_raster_draw(None, -1, pr.data_ptr(), None, 2, img.data_ptr()))


@pytest.mark.gpu
def test_render_scene_robocup(cuda_lib, oracle):
    from parallax_b200 import scene as sc
    from parallax_b200 import viz

    spec = sc.robocup_scene(n_robots=12, areas_collide=False)
    B = 16
    rng = np.random.default_rng(3)
    st = np.repeat(spec.init_state[:, :, None], B, axis=2).astype(np.float32)
    st[4:, 0:2] += rng.normal(0, 0.5, (spec.nb - 4, 2, B)).astype(np.float32)
    st = st.reshape(spec.nb * 6, B)
    frames = viz.render_scene(spec, st, env_index=[0, 3, 15]).cpu().numpy()
    assert np.array_equal(frames, oracle.raster_draw(viz.scene_primitives(spec, st, [0, 3, 15])))
    assert (frames[0] == (0, 128, 255)).all(-1).any() and (frames[0] == (128, 128, 128)).all(-1).any()

