"""Pins the oracle's restatement of jax.random (an un-vendored dependency of the reference,
pyproject.toml:23; call sites cotix/_colliders.py:142,175,222-224,254,264,284,295,
cotix/_lunar_lander.py:109-123, cotix/_collisions.py:287-298, examples/test_viz.py:39) on
published known-answer vectors: Random123's Threefry-2x32 / Philox-4x32 KATs and the outputs the
jax documentation prints for PRNGKey(0) (non-partitionable threefry, the default of every jax
release the reference can run on).  CPU only."""
import numpy as np


def test_threefry2x32_random123_kats(oracle):
    kats = [((0, 0), (0, 0), (0x6B200159, 0x99BA4EFE)),
            ((0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF, 0xFFFFFFFF), (0x1CB996FC, 0xBB002BE7)),
            ((0x13198A2E, 0x03707344), (0x243F6A88, 0x85A308D3), (0xC4923A9C, 0x483DF7A0))]
    for key, ctr, want in kats:
        out = oracle.threefry2x32(key[0], key[1], [ctr[0]], [ctr[1]])[0]
        assert (int(out[0]), int(out[1])) == want


def test_philox4x32_random123_kats(oracle):
    kats = [((0, 0), (0, 0, 0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
            ((0xFFFFFFFF, 0xFFFFFFFF), (0xFFFFFFFF,) * 4, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
            ((0xA4093822, 0x299F31D0), (0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344),
             (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1))]
    for key, ctr, want in kats:
        out = oracle.philox4x32(key[0], key[1], np.array([ctr], np.uint32))[0]
        assert tuple(int(v) for v in out) == want


def test_jax_documented_outputs(oracle):
    """Values printed in the jax documentation ("Pseudorandom numbers" / jax.random module docs,
    threefry2x32, jax_threefry_partitionable=False):
        key = PRNGKey(0)                      -> [0 0]
        random.normal(key, (1,))              -> [-0.20584226]
        key, subkey = random.split(key)       -> [4146024105 967050713], [2718843009 1272950319]
        random.normal(subkey, (1,))           -> [-1.2515389]
    """
    s = oracle.split((0, 0), 2)
    assert s.tolist() == [[4146024105, 967050713], [2718843009, 1272950319]]
    assert np.float32(oracle.normal((0, 0), 1)[0]) == np.float32(-0.20584226)
    assert np.float32(oracle.normal((2718843009, 1272950319), 1)[0]) == np.float32(-1.2515389)


def test_random_bits_layout(oracle):
    """bits for n words: counters iota(n) padded to even, first half in lane x0, second half in
    lane x1, outputs concatenated."""
    key = (123, 456)
    for n in (1, 2, 5, 8, 10):
        bits = oracle.random_bits(key, n)
        m = n + (n & 1)
        ctr = np.arange(m, dtype=np.uint32)
        ctr[n:] = 0
        h = m // 2
        out = oracle.threefry2x32(key[0], key[1], ctr[:h], ctr[h:])
        want = np.concatenate([out[:, 0], out[:, 1]])[:n]
        assert np.array_equal(bits, want)
    # split(key, m) == 2m words of bits reshaped (m, 2)
    assert np.array_equal(oracle.split(key, 5), oracle.random_bits(key, 10).reshape(5, 2))


def test_uniform_bernoulli_choice(oracle):
    key = (7, 9)
    bits = oracle.random_bits(key, 1000)
    u = oracle.uniform(key, 1000, -5.0, 5.0)
    f = ((bits >> 9) | np.uint32(0x3F800000)).view(np.float32) - np.float32(1.0)
    want = np.maximum(np.float32(-5.0), f * np.float32(10.0) + np.float32(-5.0))
    assert np.array_equal(u, want)
    assert (u >= -5).all() and (u < 5).all() and abs(u.mean()) < 0.5
    # bernoulli(key, p) = uniform(key) < p
    heads = sum(oracle.bernoulli((1, i), 0.5) for i in range(2000))
    assert 850 < heads < 1150
    for i in range(50):
        assert oracle.bernoulli((3, i), 0.5) == (oracle.uniform((3, i), 1)[0] < 0.5)
    # choice(p): only entries with p > 0 are ever drawn; roughly uniform over them
    p = np.float32([0, 1, 0, 1]) / 2
    picks = [oracle.choice((5, i), p) for i in range(400)]
    assert set(picks) == {1, 3} and 120 < picks.count(1) < 280


def test_gjk_start_direction_constant(oracle):
    """random_direction(PRNGKey(1)) (cotix/_geometry_utils.py:37-46 via _collisions.py:287-298):
    normalised jr.normal(key, (2,)).  SURVEY App. C probe: (-0.05243402, 0.9986244)."""
    P = oracle.default_lunar_params()
    d = np.float32([P.init_dir[0], P.init_dir[1]])
    nrm = oracle.normal((0, 1), 2)
    assert np.allclose(nrm, [-0.1161704, 2.2125063], atol=2e-6)
    assert np.allclose(d, nrm / np.linalg.norm(nrm), atol=1e-7)
    assert np.allclose(d, [-0.05243402, 0.9986244], atol=1e-6)
