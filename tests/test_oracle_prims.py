"""Pins the oracle's arithmetic primitives (CPU only).

 1. known-answer tests the reference's own test-suite holds for the path:
      M/tests/rand.cpp:131-141  RandomRange.UpperLimit
      M/tests/rand.cpp:19-117   range / independence properties of split streams
      M/tests/math.cpp:23-48    four golden quaternions (1e-4)
    and the published Random123 Threefry2x32-20 zero vector (initKey(0)).
 2. the restated primitives against the SAME file compiled over the reference's own rand.hpp / math.hpp
    (oracle/_ref, built only where /root/reference is mounted; otherwise the committed fixture
    tests/golden/prims_ref.npz, generated from that build by tests/golden/make_golden.py, is used).
"""
import ctypes
import os

import numpy as np
import pytest

from oracle_binding import oracle_lib

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
u32 = ctypes.c_uint32


def _split(lib, a, b, idx, up=0):
    out = (u32 * 2)()
    lib.mbo_prim_split(u32(a), u32(b), u32(idx), u32(up), out)
    return out[0], out[1]


def _init_key(lib, seed):
    out = (u32 * 2)()
    lib.mbo_prim_init_key(u32(seed), out)
    return out[0], out[1]


@pytest.fixture(scope="module")
def lib():
    return oracle_lib(False)


@pytest.fixture(scope="module")
def ref():
    return oracle_lib(True)   # None where the reference is not mounted and oracle/_ref was not shipped


def test_flavours(lib, ref):
    assert lib.mbo_flavour() == b"restated"
    if ref is not None:
        assert ref.mbo_flavour() == b"reference-headers"


def test_threefry_zero_vector(lib):
    # Random123 kat_vectors: threefry2x32 20 rounds, counter 0 0, key 0 0 -> 6b200159 99ba4efe
    assert _init_key(lib, 0) == (0x6B200159, 0x99BA4EFE)
    # world 0 of seed 0 (src/sim/sim.cpp:1238-1239): split_i(initKey(0), 0, 0), probed against the
    # reference headers in SURVEY.md section 8c
    assert _split(lib, 0x6B200159, 0x99BA4EFE, 0, 0) == (0xF84E8312, 0x2FEF64F3)


def test_threefry_random123_vectors(lib):
    # the two other published threefry2x32_20 known answers (counter, key all-ones / digits of pi)
    assert _split(lib, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF) == (0x1CB996FC, 0xBB002BE7)
    assert _split(lib, 0x13198A2E, 0x03707344, 0x243F6A88, 0x85A308D3) == (0xC4923A9C, 0x483DF7A0)


def test_rand_upper_limit_kat(lib):
    # M/tests/rand.cpp:131-141: key {0xFFFFFFFF, 0}: bits32 == 0xFFFFFFFF, sampleI32(k, 0, 64) == 63
    assert lib.mbo_prim_sample_i32(u32(0xFFFFFFFF), u32(0), 0, 64) == 63
    # sampleUniform of the same key is the largest float below 1
    assert lib.mbo_prim_sample_uniform(u32(0xFFFFFFFF), u32(0)) == np.float32(1.0 - 2.0 ** -24)


@pytest.mark.parametrize("lo,hi", [(0, 5), (2, 20), (-20, 2), (-30, -10), (-2 ** 31, -2 ** 31 + 20)])
def test_rand_split_stream_limits(lib, lo, hi):
    # M/tests/rand.cpp:19-117 (RandomSplitTest): two streams split from RandKey(5) stay in range
    # and do not collide; RNG(key) draws are split_i(key, count++) (rand.inl:249-255)
    ka, kb = _split(lib, 5, 0, 0), _split(lib, 5, 0, 1)
    same = 0
    for i in range(100):
        da, db = _split(lib, *ka, i), _split(lib, *kb, i)
        va = lib.mbo_prim_sample_i32(u32(da[0]), u32(da[1]), lo, hi)
        vb = lib.mbo_prim_sample_i32(u32(db[0]), u32(db[1]), lo, hi)
        assert lo <= va < hi and lo <= vb < hi
        ua = lib.mbo_prim_sample_uniform(u32(da[0]), u32(da[1]))
        ub = lib.mbo_prim_sample_uniform(u32(db[0]), u32(db[1]))
        assert 0.0 <= ua < 1.0 and 0.0 <= ub < 1.0
        same += ua == ub
    assert same == 0


def test_rand_different_seeds(lib):
    # M/tests/rand.cpp:119-129
    ka, kb = _init_key(lib, 5), _init_key(lib, 10)
    for i in range(100):
        da, db = _split(lib, *ka, i), _split(lib, *kb, i)
        assert lib.mbo_prim_sample_uniform(u32(da[0]), u32(da[1])) != lib.mbo_prim_sample_uniform(u32(db[0]), u32(db[1]))


def test_sample_i32_unbiased_small_range(lib):
    # Lemire rejection (rand.inl:115-159): exact uniformity over a range that does not divide 2^32
    rng = np.random.default_rng(0)
    keys = rng.integers(0, 2 ** 32, size=(30000, 2), dtype=np.uint64)
    hist = np.zeros(10, dtype=np.int64)
    for a, b in keys:
        hist[lib.mbo_prim_sample_i32(u32(int(a)), u32(int(b)), 0, 10)] += 1
    assert hist.min() > 2700 and hist.max() < 3300, hist


def _quat(lib, name, *args):
    out = (ctypes.c_float * 4)()
    getattr(lib, name)(*args, out)
    return np.array(out[:], dtype=np.float32)


def test_math_quaternion_kats(lib, ref):
    # M/tests/math.cpp:23-48. The oracle restates only rotations about z (the simulator never uses another
    # axis), so the y- and x-axis cases of the reference test are checked through the product formula:
    # q(45deg about y) * q(45deg about x) == (0.853553, 0.353553, 0.353553, -0.146447)
    f4 = ctypes.c_float * 4
    lib.mbo_prim_quat_mul.argtypes = [f4, f4, f4]
    q2 = f4(0.9238795, 0.0, 0.3826834, 0.0)
    q3 = f4(0.9238795, 0.3826834, 0.0, 0.0)
    out = f4()
    lib.mbo_prim_quat_mul(q2, q3, out)
    np.testing.assert_allclose(out[:], [0.853553, 0.353553, 0.353553, -0.146447], atol=1e-4)
    # angleAxis(0, axis) == identity; angleAxis(45deg, z) == (cos 22.5deg, 0, 0, sin 22.5deg)
    np.testing.assert_allclose(_quat(lib, "mbo_prim_angle_axis_z", ctypes.c_float(0.0)), [1, 0, 0, 0], atol=1e-4)
    np.testing.assert_allclose(_quat(lib, "mbo_prim_angle_axis_z", ctypes.c_float(np.pi / 4)),
                               [0.9238795, 0, 0, 0.3826834], atol=1e-4)
    # SURVEY.md 8c probe of the reference headers: angleAxis(0.1, z) * x-hat = (0.995004177, 0.099833414, 0)
    q = _quat(lib, "mbo_prim_angle_axis_z", ctypes.c_float(0.1))
    f3 = ctypes.c_float * 3
    lib.mbo_prim_rotate.argtypes = [f4, f3, f3]
    v = f3()
    lib.mbo_prim_rotate(f4(*q), f3(1, 0, 0), v)
    np.testing.assert_allclose(v[:], [0.995004177, 0.099833414, 0.0], atol=1e-7)


def test_turn_constants_are_correctly_rounded(lib):
    # oracle_prims.h hard-codes cosf(0.05f) / sinf(0.05f); both must be the float nearest the real value
    q = _quat(lib, "mbo_prim_angle_axis_z", ctypes.c_float(0.1))
    half = np.float64(np.float32(0.1) / np.float32(2.0))
    assert q[0] == np.float32(np.cos(half)) and q[3] == np.float32(np.sin(half))
    qn = _quat(lib, "mbo_prim_angle_axis_z", ctypes.c_float(-0.1))
    assert qn[0] == q[0] and qn[3] == -q[3] and qn[1] == 0 and qn[2] == 0


def _prim_table(lib, seed=123, n=4000):
    """Deterministic table of primitive outputs (the fixture compared across flavours)."""
    rng = np.random.default_rng(seed)
    keys = rng.integers(0, 2 ** 32, size=(n, 4), dtype=np.uint64)
    split = np.array([_split(lib, int(a), int(b), int(c), int(d) & 0xFFFF) for a, b, c, d in keys], dtype=np.uint32)
    ints = np.array([lib.mbo_prim_sample_i32(u32(int(a)), u32(int(b)), -7, 23) for a, b, _, _ in keys], dtype=np.int32)
    unif = np.array([lib.mbo_prim_sample_uniform(u32(int(a)), u32(int(b))) for a, b, _, _ in keys], dtype=np.float32)
    # a random walk of +-0.1 rad turns, the rotated x axis and its normalisation after every turn
    f4, f3 = ctypes.c_float * 4, ctypes.c_float * 3
    lib.mbo_prim_quat_mul.argtypes = [f4, f4, f4]
    lib.mbo_prim_rotate.argtypes = [f4, f3, f3]
    lib.mbo_prim_normalize.argtypes = [f3, f3]
    left = f4(*_quat(lib, "mbo_prim_angle_axis_z", ctypes.c_float(0.1)))
    right = f4(*_quat(lib, "mbo_prim_angle_axis_z", ctypes.c_float(-0.1)))
    q = f4(1, 0, 0, 0)
    walk = np.zeros((1500, 10), dtype=np.float32)
    turns = rng.integers(0, 2, size=1500)
    for i, t in enumerate(turns):
        o = f4()
        lib.mbo_prim_quat_mul(q, left if t else right, o)
        q = o
        v, nv = f3(), f3()
        lib.mbo_prim_rotate(q, f3(1, 0, 0), v)
        lib.mbo_prim_normalize(v, nv)
        walk[i] = list(q) + list(v) + list(nv)
    return {"split": split, "ints": ints, "unif": unif, "walk": walk}


def test_restated_prims_equal_reference_headers(lib, ref):
    """Bit-for-bit: restated Threefry / Lemire / quaternion formulas == the reference's rand.hpp / math.hpp."""
    mine = _prim_table(lib)
    if ref is not None:
        theirs = _prim_table(ref)
        src = "oracle/_ref"
    else:
        path = os.path.join(GOLDEN, "prims_ref.npz")
        assert os.path.exists(path), "neither oracle/_ref nor tests/golden/prims_ref.npz is available"
        theirs = dict(np.load(path))
        src = "tests/golden/prims_ref.npz"
    for k in mine:
        a, b = mine[k], theirs[k]
        assert a.dtype == b.dtype and a.shape == b.shape, (k, src)
        np.testing.assert_array_equal(a.view(np.uint32) if a.dtype == np.float32 else a,
                                      b.view(np.uint32) if b.dtype == np.float32 else b, err_msg=f"{k} vs {src}")


def test_golden_prims_fixture_matches_ref_build(ref):
    """Where the reference is mounted, the committed fixture must be what its headers produce today."""
    path = os.path.join(GOLDEN, "prims_ref.npz")
    if ref is None or not os.path.exists(path):
        pytest.skip("needs both /root/reference and the committed fixture")
    gold = dict(np.load(path))
    now = _prim_table(ref)
    for k in now:
        np.testing.assert_array_equal(now[k].view(np.uint32) if now[k].dtype == np.float32 else now[k],
                                      gold[k].view(np.uint32) if gold[k].dtype == np.float32 else gold[k], err_msg=k)
