"""The product's arithmetic header (madrona-bots_b200/csrc/sim_prims.cuh) compiled for the host and compared
bit-for-bit with the oracle's primitives, plus the proof-by-test that the culled sensor of export_kernel equals
the exhaustive sensor (CPU only; the same header is what nvcc compiles into the kernels)."""
import ctypes

import numpy as np
import pytest

from oracle_binding import oracle_lib
from prims_binding import shim

u32, f32 = ctypes.c_uint32, ctypes.c_float


@pytest.fixture(scope="module")
def S():
    return shim()


@pytest.fixture(scope="module")
def O():
    lib = oracle_lib(False)
    f4, f3 = f32 * 4, f32 * 3
    lib.mbo_prim_quat_mul.argtypes = [f4, f4, f4]
    lib.mbo_prim_rotate.argtypes = [f4, f3, f3]
    lib.mbo_prim_normalize.argtypes = [f3, f3]
    return lib


def bits(x):
    return np.float32(x).view(np.uint32)


def test_threefry_and_samplers_match_oracle(S, O):
    rng = np.random.default_rng(5)
    out_s, out_o = (u32 * 2)(), (u32 * 2)()
    for a, b, c, d in rng.integers(0, 2 ** 32, size=(3000, 4), dtype=np.uint64):
        a, b, c, d = int(a), int(b), int(c), int(d)
        S.shim_split(a, b, c, d, out_s)
        O.mbo_prim_split(u32(a), u32(b), u32(c), u32(d), out_o)
        assert out_s[:] == out_o[:]
        for lo, hi in ((0, 10), (1, 3), (0, 8), (0, 6), (0, 16), (-5, 7)):
            assert S.shim_sample_i32(a, b, lo, hi) == O.mbo_prim_sample_i32(u32(a), u32(b), lo, hi)
        assert bits(S.shim_sample_uniform(a, b)) == bits(O.mbo_prim_sample_uniform(u32(a), u32(b)))
    S.shim_init_key(0, out_s)
    assert out_s[:] == [0x6B200159, 0x99BA4EFE]


def test_planar_rotation_equals_full_quaternion(S, O):
    """(w, z) pair under rot_turn / rot_forward / dir_normalize == Quat * angleAxis, rotateVec(x-hat), normalize of the
    oracle, bit for bit, along a long random walk of turns (the planar formulas only drop +-0 terms)."""
    f4, f3 = f32 * 4, f32 * 3
    left, right = f4(), f4()
    O.mbo_prim_angle_axis_z(f32(0.1), left)
    O.mbo_prim_angle_axis_z(f32(-0.1), right)
    q = f4(1, 0, 0, 0)
    w, z = np.float32(1), np.float32(0)
    rng = np.random.default_rng(9)
    o2 = (f32 * 2)()
    for t in rng.integers(0, 2, size=4000):
        nq = f4()
        O.mbo_prim_quat_mul(q, left if t else right, nq)
        q = nq
        S.shim_turn(w, z, int(t), o2)
        w, z = np.float32(o2[0]), np.float32(o2[1])
        assert bits(w) == bits(q[0]) and bits(z) == bits(q[3]) and q[1] == 0 and q[2] == 0
        v, nv = f3(), f3()
        O.mbo_prim_rotate(q, f3(1, 0, 0), v)
        O.mbo_prim_normalize(v, nv)
        S.shim_forward(w, z, o2)
        assert bits(o2[0]) == bits(v[0]) and bits(o2[1]) == bits(v[1]) and v[2] == 0
        S.shim_normalize(v[0], v[1], o2)
        assert bits(o2[0]) == bits(nv[0]) and bits(o2[1]) == bits(nv[1])


def test_move_clamps_and_precedence(S):
    """actionSystem movement (src/sim/sim.cpp:456-486): left beats right, forward beats backward, clamp to
    [0, 127] x [0, 95]."""
    def move(x, y, w, z, f, b, l, r):
        cx, cy, cw, cz = f32(x), f32(y), f32(w), f32(z)
        ln = S.shim_move(ctypes.byref(cx), ctypes.byref(cy), ctypes.byref(cw), ctypes.byref(cz), f, b, l, r)
        return cx.value, cy.value, cw.value, cz.value, ln
    x, y, w, z, ln = move(10, 10, 1, 0, 1, 1, 0, 0)
    assert (x, y) == (11.0, 10.0) and ln == 1.0
    x, y, w, z, ln = move(10, 10, 1, 0, 0, 1, 0, 0)
    assert (x, y) == (9.0, 10.0)
    _, _, w1, z1, _ = move(10, 10, 1, 0, 0, 0, 1, 1)
    _, _, w2, z2, _ = move(10, 10, 1, 0, 0, 0, 1, 0)
    assert (w1, z1) == (w2, z2) and z1 > 0
    x, y, _, _, ln = move(126.5, 0.2, 1, 0, 1, 0, 0, 0)
    assert x == 127.0 and ln == 0.5
    x, y, _, _, ln = move(0.3, 5, 1, 0, 0, 1, 0, 0)
    assert x == 0.0
    # facing +y: rotation by 90 degrees
    c = np.float32(np.cos(np.pi / 4)); s = np.float32(np.sin(np.pi / 4))
    x, y, _, _, _ = move(5, 94.9, c, s, 1, 0, 0, 0)
    assert y == 95.0


def test_chunk_cell_and_surroundings(S):
    rng = np.random.default_rng(2)
    na = rng.integers(0, 9, size=48).astype(np.uint32)
    ts = rng.integers(0, 30, size=48).astype(np.uint32)
    pa = na.ctypes.data_as(ctypes.POINTER(u32)); pt = ts.ctypes.data_as(ctypes.POINTER(u32))
    out = (f32 * 2)()
    for _ in range(3000):
        x, y = np.float32(rng.random() * 127), np.float32(rng.random() * 95)
        c = S.shim_chunk_of(x, y)
        assert c == int(x // 16) + int(y // 16) * 8
        cell = S.shim_cell_of(x, y)
        assert cell >> 4 == int(np.float32(16) * (np.float32(x * np.float32(0.0625)) - np.floor(np.float32(x * np.float32(0.0625)))))
        S.shim_surroundings(x, y, pa, pt, out)
        # float64 restatement of updateSurroundingObservation (src/sim/sim.cpp:583-654)
        chx, chy = (float(x) - 8) / 16, (float(y) - 8) / 16
        fx, fy, cx, cy = np.floor(chx), np.floor(chy), np.ceil(chx), np.ceil(chy)

        def at(arr, ix, iy):
            return float(arr[int(ix) + int(iy) * 8]) if 0 <= ix < 8 and 0 <= iy < 6 else 0.0
        xi, yi = chx - fx, chy - fy
        for k, arr in enumerate((na, ts)):
            v0 = xi * at(arr, cx, fy) + (1 - xi) * at(arr, fx, fy)
            v1 = xi * at(arr, cx, cy) + (1 - xi) * at(arr, fx, cy)
            assert abs(out[k] - (yi * v1 + (1 - yi) * v0)) < 1e-4
    # corners and edges
    for x, y in ((0, 0), (127, 95), (0, 95), (127, 0), (16, 16), (8, 8)):
        assert 0 <= S.shim_chunk_of(x, y) < 48
    assert S.shim_chunk_of(127, 95) == 47 and S.shim_cell_of(127, 95) == (15 << 4 | 15)


def test_rewards(S):
    # speciesInfoSync (src/sim/sim.cpp:809-818) and rewardSystem setting 8 (:943-956)
    assert S.shim_species_reward(0, 0, 32) == -2.0
    assert S.shim_species_reward(8, 800, 32) == np.float32(np.float32(0.25) + np.float32(1.0) - np.float32(2.0))
    r = S.shim_agent_reward(f32(-0.75), 70, 1, 0, 1)
    exp = np.float32(np.float32(np.float32(-0.75) + np.float32(70) / np.float32(100)) - np.float32(0.5)) + np.float32(10) + np.float32(15)
    assert r == exp


@pytest.mark.parametrize("mode,n,trials", [(0, 33, 400), (0, 128, 40), (1, 40, 300), (2, 24, 3000)])
def test_culled_sensor_equals_exhaustive(S, mode, n, trials):
    """pair_candidates + the exact test on candidates only == the exhaustive 32-ray sensor, on random worlds,
    crowded patches and targets placed within ulps of a ray's tangent."""
    tests = ctypes.c_longlong()
    bad = S.shim_cull_trials(1234 + mode, trials, n, mode, ctypes.byref(tests))
    assert bad == 0
    observers = trials * (1 if mode == 2 else n)
    per_pair = tests.value / (observers * n)
    assert per_pair < 32          # the cull must actually cull
    if mode == 0 and n == 33:
        assert per_pair < 0.6, per_pair


def test_wedge_stage_never_rejects_a_pair_with_candidates(S):
    """Stage 1 of the cull (pair_maybe) is conservative w.r.t. stage 2 (pair_run), and rejects about half of
    uniformly random pairs."""
    passed = ctypes.c_longlong()
    n = 4_000_000
    assert S.shim_wedge_trials(77, n, ctypes.byref(passed)) == 0
    assert passed.value < 0.85 * n
