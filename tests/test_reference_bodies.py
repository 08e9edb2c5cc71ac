"""The CPU oracle pinned to the REFERENCE'S OWN SOURCE TEXT (VERDICT r1 "missing" #3).

oracle/_ref/librefsim.so (recipe: `make -C oracle refsim`; sources: oracle/refmock/) compiles, from where they lie
under /root/reference, src/sim/sim.cpp (every system body, Sim::Sim, the task graphs), external/madrona/src/render/
ecs_system.cpp (view / instance transforms) and raySphereIntersect / traceRayTLAS / bvhRaycastEntry of
external/madrona/src/mw/device/bvh_raycast.cpp, and runs ONE world through the reference's own task graphs under a
host mock of the ECS runtime. These tests drive that world and the oracle with the same actions:

  * with the rules of DEVIATIONS.md switched on (U5, U6, U8, U19, D3; D2 at compile time) every export of the
    oracle equals the reference text's, bit for bit, every step -- positions, rotations, health, food, RNG stream,
    rewards, surroundings, statistics, 32-ray depth rows, finder targets, prev columns;
  * with one rule switched off at a time the LITERAL text differs from the oracle exactly as DEVIATIONS.md says.

Where /root/reference is absent (the GPU box) the live comparison needs the shipped .so; the committed fixture
tests/golden/refsim_rollouts.npz (generated from the same library) is checked either way.
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))

from oracle_binding import OracleSim, random_actions  # noqa: E402
import refsim_binding as rb  # noqa: E402
from make_refsim_golden import CASES, digest, food_canonical, run_case  # noqa: E402

needs_refsim = pytest.mark.skipif(rb.refsim_lib() is None, reason="oracle/_ref/librefsim.so not built (needs /root/reference)")


class _OracleOneWorld(OracleSim):
    def __init__(self, seed, world_id, init):
        super().__init__(1, seed, init, world_id_offset=world_id)

    def write(self, name, value):
        if len(value):
            self.tensor(name)[:] = value


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_reproduces_reference_text_golden(name):
    """No reference needed at run time: the oracle against digests generated from the reference text."""
    gold = np.load(os.path.join(HERE, "golden", "refsim_rollouts.npz"))
    sim, counts, digs = run_case(_OracleOneWorld, CASES[name])
    np.testing.assert_array_equal(counts, gold[name + ".counts"])
    bad = np.argwhere(digs != gold[name + ".digests"])
    assert len(bad) == 0, f"first mismatch: step {bad[0][0]}, export {rb.REF_EXPORTS[bad[0][1]]}"
    if name == "s0_w0":
        for n in rb.REF_EXPORTS:
            want = gold["s0_w0.final." + n]
            have = food_canonical(sim.tensor(n)) if n == "DEBUG_FOOD" else sim.tensor(n)
            np.testing.assert_array_equal(have, food_canonical(want) if n == "DEBUG_FOOD" else want, err_msg=n)


def _drive_pair(seed, wid, init, steps, mode, flags=rb.CANONICAL, libm=False, on_step=None):
    r = rb.RefSimAdapter(seed, wid, init, flags, libm)
    o = _OracleOneWorld(seed, wid, init)
    rng = np.random.default_rng(seed * 7 + wid)
    for t in range(steps):
        A = o.num_agents()
        assert r.num_agents() == A
        acts = random_actions(rng, A, mode)
        hid = rng.standard_normal((A, 16)).astype(np.float32)
        for sim in (o, r):
            sim.write("ACTION", acts)
            sim.write("HIDDEN_STATE", hid)
        o.step(); r.step()
        if on_step:
            on_step(t, o, r)
        if t % 2 == 0:
            o.shift_observations(); r.shift_observations()
    return o, r


def _assert_all_equal(o, r, where, skip=()):
    for n in rb.REF_EXPORTS:
        if n in skip:
            continue
        a, b = o.tensor(n), r.tensor(n)
        if n == "DEBUG_FOOD":
            a, b = food_canonical(a), food_canonical(b)
        assert a.shape == b.shape, f"{where}: {n} shape {a.shape} != reference text {b.shape}"
        np.testing.assert_array_equal(a, b, err_msg=f"{where}: {n}")


@needs_refsim
@pytest.mark.parametrize("seed,wid,init,steps,mode", [(0, 0, 32, 120, "bits"), (3, 7, 32, 150, "onehot"),
                                                      (11, 123456, 32, 100, "bits"), (5, 2, 7, 80, "bits"),
                                                      (8, 1, 60, 40, "bits")])
def test_oracle_systems_equal_reference_bodies(seed, wid, init, steps, mode):
    """Every export, every step, against the reference's own system bodies run in the reference's own node order."""
    stats = {}

    def check(t, o, r):
        _assert_all_equal(o, r, f"seed {seed} world {wid} step {t}")
        stats["st"] = o.episode_stats()

    _drive_pair(seed, wid, init, steps, mode, on_step=check)
    st = stats["st"]
    if init >= 32:
        assert st[10] > 0 and st[11] > 0 and st[12] > 0, f"the rollout must exercise births, deaths and respawns: {st}"


@needs_refsim
def test_free_functions_of_the_reference():
    """raySphereIntersect (bvh_raycast.cpp:70-92) against the oracle's arithmetic on random and near-tangent rays."""
    import ctypes
    from prims_binding import shim
    lib = rb.refsim_lib()
    rng = np.random.default_rng(1)
    L = shim()
    have_shim = hasattr(L, "shim_ray_sphere")
    n_hit = 0
    for k in range(4000):
        o = rng.uniform(0, 128, 3).astype(np.float32); o[2] = 1.0
        ang = rng.uniform(-np.pi, np.pi)
        d = np.array([np.cos(ang), np.sin(ang), 1e-5], dtype=np.float32)
        dist = rng.uniform(1.0, 60.0)
        perp = rng.uniform(-1.2, 1.2) if k % 2 else rng.choice([-1.0, 1.0]) * (1.0 + rng.uniform(-1e-4, 1e-4))
        c = (o + dist * d + perp * np.array([-d[1], d[0], 0], dtype=np.float32)).astype(np.float32); c[2] = 1.0
        t = ctypes.c_float()
        hit = lib.mbr_ray_sphere(c.ctypes.data, ctypes.c_float(1.0), o.ctypes.data, d.ctypes.data, ctypes.byref(t))
        # the oracle's formula, float32 step by step (oracle/oracle_prims.h ray_sphere == sim_prims.cuh ray_sphere)
        f = np.float32
        oc = (o - c).astype(np.float32)
        a = f(f(f(d[0] * d[0]) + f(d[1] * d[1])) + f(d[2] * d[2]))
        b = f(f(2.0) * f(f(f(oc[0] * d[0]) + f(oc[1] * d[1])) + f(oc[2] * d[2])))
        cc = f(f(f(f(oc[0] * oc[0]) + f(oc[1] * oc[1])) + f(oc[2] * oc[2])) - f(1.0))
        disc = f(f(b * b) - f(f(f(4.0) * a) * cc))
        assert bool(hit) == bool(disc >= 0), (k, disc)
        if hit:
            n_hit += 1
            want = f(f(-b - np.sqrt(disc, dtype=np.float32)) / f(f(2.0) * a))
            assert t.value == want, (k, t.value, want)
    assert n_hit > 1000


# ---- the literal text, one rule off at a time ----------------------------------------------------------------
@needs_refsim
def test_literal_U5_reward_reads_the_next_species_term():
    """rewardSystem indexes rewards[speciesID] with the 1-based id (sim.cpp:943): literally, species s gets the term
    of species s + 1 and species 4 reads past the row. Everything else is unaffected."""
    seen = {"diff": 0}

    def check(t, o, r):
        _assert_all_equal(o, r, f"U5 step {t}", skip=("REWARD", "PREV_REWARD"))
        oi, of, food, chunks = r.w.world()                         # SpeciesReward of this step (0-based, as written)
        f32 = np.float32
        species = r.tensor("DEBUG_AGENT_SPECIES")[:, 0]
        rows = r.tensor("SENSOR_INDEX")[:, 0]
        health = r.tensor("DEBUG_AGENT_HEALTH")[:, 0]
        stats = o.tensor("STATS")                                   # row order; (hitFriendly, hitEnemy, ate, reproduced)
        lit = r.tensor("REWARD")[:, 0]
        can = o.tensor("REWARD")[:, 0]
        for a in range(len(species)):
            s, row = int(species[a]), int(rows[a])
            term_lit = of[s] if s < 4 else f32(0.0)                 # what follows the row in the mock is zero
            def reward(term):
                v = f32(f32(term + f32(f32(health[a]) / f32(100.0))) - f32(0.5))
                if stats[row, 2]: v = f32(v + f32(10.0))
                if stats[row, 3]: v = f32(v + f32(10.0))
                if stats[row, 1]: v = f32(v + f32(15.0))
                return v
            assert lit[row] == reward(term_lit), (t, a, s)
            assert can[row] == reward(of[s - 1]), (t, a, s)
            seen["diff"] += int(lit[row] != can[row])

    _drive_pair(2, 5, 32, 40, "bits", flags=rb.CANONICAL & ~rb.FIX_U5, on_step=check)
    assert seen["diff"] > 0


@needs_refsim
def test_literal_U6_observation_rows_leak():
    """Dying agents destroy only the renderable (sim.cpp:574-578): literally the observation table keeps one stale
    row per death while the Agent table -- the tensors' leading dimension -- matches the oracle."""
    # driven without actions: a literal observation row has no counterpart in the oracle to take its action from
    o2 = _OracleOneWorld(4, 9, 32)
    r2 = rb.RefSimAdapter(4, 9, 32, rb.CANONICAL & ~rb.FIX_U6)
    dead_total = 0
    for t in range(200):
        o2.step(); r2.step()
        dead_total = int(o2.episode_stats()[11])
        assert r2.num_agents() == o2.num_agents()
        assert r2.w.num_obs_rows() == o2.num_agents() + dead_total, (t, r2.w.num_obs_rows(), o2.num_agents(), dead_total)
        np.testing.assert_array_equal(r2.tensor("DEBUG_POSE"), o2.tensor("DEBUG_POSE"))
        np.testing.assert_array_equal(r2.tensor("DEBUG_AGENT_HEALTH"), o2.tensor("DEBUG_AGENT_HEALTH"))


@needs_refsim
def test_literal_U8_traversal_order_hits_are_never_nearer():
    """traceRayTLAS keeps the last accepted hit of its traversal, not the nearest (bvh_raycast.cpp:173-180): literally a
    few rays report a farther agent. Never a nearer one, never a hit where the nearest-hit rule has a miss."""
    tot = {"rays": 0, "diff": 0}

    def check(t, o, r):
        a, b = o.tensor("SENSOR_DEPTH")[:, 1:].astype(int), r.tensor("SENSOR_DEPTH")[:, 1:].astype(int)
        assert (b >= a).all(), f"step {t}: the literal traversal reported a nearer hit than the nearest"
        assert ((a == 255) == (b == 255)).all()
        tot["rays"] += a.size
        tot["diff"] += int((a != b).sum())

    # one step at a time from identical states: feed both sides the oracle's state by restarting pairs
    for seed in range(6):
        _drive_pair(seed, seed + 1, 32, 1, "onehot", flags=rb.CANONICAL & ~rb.FIX_U8, on_step=check)
    assert tot["diff"] <= tot["rays"] // 50, tot


@needs_refsim
def test_literal_U19_sensor_rows_follow_the_agent_table():
    """viewTransformUpdate resets rowIDX to the Agent-table row (ecs_system.cpp:184): literally the pixel-0 stamp of the
    depth row written for an agent is its Agent-table row, not its observation row."""
    r = rb.RefSimAdapter(1, 0, 32, rb.CANONICAL & ~rb.FIX_U19)
    r.step()
    depth = r.tensor("SENSOR_DEPTH")
    # row k of the sensor buffer was written by the agent whose AGENT row is k
    assert (depth[:r.num_agents(), 0] == np.arange(r.num_agents(), dtype=np.uint8)).all()
    c = rb.RefSimAdapter(1, 0, 32)
    c.step()
    rows = c.tensor("SENSOR_INDEX")[:, 0]
    assert (c.tensor("SENSOR_DEPTH")[rows, 0] == (rows & 0xFF).astype(np.uint8)).all()
    assert not (rows == np.arange(len(rows))).all()


@needs_refsim
def test_rule_D2_libm_differs_only_in_rare_sensor_bytes():
    """The same text with the host's libm cosf / sinf / atan2f instead of include/mbots_detmath.h: positions, health and
    the RNG stream do not depend on them; depth bytes differ for a ray in ~10^4 (a hit distance within an ulp of an
    integer, a grazing ray). Measured from identical states (first step of several worlds)."""
    rays = diff = 0
    for seed in range(8):
        a = rb.RefSimAdapter(seed, 3, 32)
        b = rb.RefSimAdapter(seed, 3, 32, libm=True)
        rng = np.random.default_rng(seed)
        acts = random_actions(rng, 32, "onehot")
        for s in (a, b):
            s.write("ACTION", acts)
            s.step()
        np.testing.assert_array_equal(a.tensor("DEBUG_POSE"), b.tensor("DEBUG_POSE"))
        np.testing.assert_array_equal(a.tensor("DEBUG_RNG_COUNT"), b.tensor("DEBUG_RNG_COUNT"))
        da, db = a.tensor("SENSOR_DEPTH"), b.tensor("SENSOR_DEPTH")
        rays += da.size
        diff += int((da != db).sum())
    assert diff <= max(2, rays // 2000), (diff, rays)
