"""The CPU oracle of the step (CPU only): golden rollouts, flavour equivalence, and the canonical-order rules
of SURVEY.md section 8a the restatement commits to."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))

from make_golden import DIGEST_EXPORTS, ROLLOUTS, OracleAdapter, run_rollout  # noqa: E402
from oracle_binding import OracleSim, oracle_lib, random_actions  # noqa: E402

GOLD = os.path.join(HERE, "golden", "rollout_ref.npz")


@pytest.fixture(scope="module")
def gold():
    return dict(np.load(GOLD))


@pytest.mark.parametrize("name", sorted(ROLLOUTS))
def test_restated_oracle_reproduces_golden_rollouts(gold, name):
    """The travelling oracle (restated primitives) == the fixtures generated through the reference's own
    rand.hpp / math.hpp: agent counts and a CRC of every export after every step."""
    sim, counts, digs = run_rollout(lambda w, s, i, c, o: OracleAdapter(w, s, i, cap=c, world_id_offset=o), ROLLOUTS[name])
    np.testing.assert_array_equal(counts, gold[name + ".counts"])
    bad = np.argwhere(digs != gold[name + ".digests"])
    assert len(bad) == 0, f"first mismatch: step {bad[0][0]}, export {DIGEST_EXPORTS[bad[0][1]]}"
    np.testing.assert_array_equal(sim.episode_stats(), gold[name + ".stats"])
    if name == "w1":
        for n in DIGEST_EXPORTS:
            np.testing.assert_array_equal(np.array(sim.tensor(n)), gold["w1.final." + n], err_msg=n)


def test_ref_flavour_reproduces_golden_rollouts(gold):
    """Where /root/reference is mounted: the fixture is what the reference-header build produces today."""
    if oracle_lib(True) is None:
        pytest.skip("oracle/_ref unavailable (reference not mounted)")
    name = "w24_bits"
    _, counts, digs = run_rollout(lambda w, s, i, c, o: OracleAdapter(w, s, i, cap=c, world_id_offset=o, ref=True), ROLLOUTS[name])
    np.testing.assert_array_equal(counts, gold[name + ".counts"])
    np.testing.assert_array_equal(digs, gold[name + ".digests"])


def test_rollouts_exercise_the_ragged_paths(gold):
    s = gold["w24_bits.stats"]
    assert s[10] > 0 and s[11] > 0 and s[12] > 0 and s[14] > 0 and s[15] > 0, s   # births deaths respawns eaten shots
    assert gold["w8_cap32.stats"][13] > 0                                            # overflow (capacity) path
    assert len(set(gold["w24_bits.counts"].tolist())) > 5


def _roll(o, steps, seed, mode="bits", shift=True):
    rng = np.random.default_rng(seed)
    for _ in range(steps):
        A = o.num_agents()
        if A:
            o.tensor("ACTION")[:] = random_actions(rng, A, mode)
            o.tensor("HIDDEN_STATE")[:] = rng.standard_normal((A, 16)).astype(np.float32)
        o.step()
        if shift:
            o.shift_observations()


def test_planar_restatement_is_exact():
    """sim_prims.cuh keeps a rotation as (w, z): the full quaternion's x, y and pos.z - 1 stay exactly 0."""
    o = OracleSim(16, 2, 32)
    _roll(o, 150, 1)
    assert o.planarity() == 0.0


def test_rows_are_species_major_and_world_ordered():
    """Canonical export order (species, world, slot) -- U11; species blocks are what
    learn/training_loop.py:43-45 slices out of every tensor."""
    o = OracleSim(20, 5, 32)
    _roll(o, 40, 3)
    rows = o.tensor("SENSOR_INDEX")[:, 0]
    species = np.array(o.tensor("DEBUG_AGENT_SPECIES")[:, 0])   # debug dumps share one scratch buffer: copy
    wc = o.tensor("DEBUG_WORLD_COUNT")[:, 0]
    world = np.repeat(np.arange(20), wc)
    assert np.array_equal(np.sort(rows), np.arange(o.num_agents()))
    order = np.argsort(rows)
    key = species[order].astype(np.int64) * 1000 + world[order]
    assert np.all(np.diff(key) >= 0)
    # position export row r holds the pose of the agent whose sensor index is r
    np.testing.assert_array_equal(o.tensor("POSITION")[rows], o.tensor("DEBUG_POSE")[:, :2])
    np.testing.assert_array_equal(o.tensor("HEALTH")[rows, 0], o.tensor("DEBUG_AGENT_HEALTH")[:, 0].astype(np.float32))
    # U9: pixel 0 of the depth row carries (uint8) of its own row index
    np.testing.assert_array_equal(o.tensor("SENSOR_DEPTH")[:, 0], (np.arange(o.num_agents()) & 0xFF).astype(np.uint8))
    # row bases: exclusive scan in (species, world) order
    rb = o.tensor("DEBUG_ROW_BASE")
    cnt = np.zeros((20, 4), dtype=np.int64)
    np.add.at(cnt, (world, species - 1), 1)
    np.testing.assert_array_equal(rb.T.reshape(-1), np.concatenate([[0], np.cumsum(cnt.T.reshape(-1))[:-1]]))


def test_shift_semantics_and_u7():
    """shiftObservationsSystem (src/sim/sim.cpp:1002-1037): prev = cur, with prev.hitEnemy = cur.hitFriendly (:1034)."""
    o = OracleSim(12, 8, 32)
    _roll(o, 30, 4, shift=False)
    cur = {n: np.array(o.tensor(n)) for n in ("POSITION", "HEALTH", "SURROUNDING", "REWARD", "ACTION", "HIDDEN_STATE", "STATS")}
    sem, dep = np.array(o.tensor("PREV_SENSOR_SEMANTIC")), np.array(o.tensor("PREV_SENSOR_DEPTH"))
    o.shift_observations()
    for n in ("POSITION", "HEALTH", "SURROUNDING", "REWARD", "ACTION", "HIDDEN_STATE"):
        np.testing.assert_array_equal(o.tensor("PREV_" + n), cur[n], err_msg=n)
    ps, st = o.tensor("PREV_STATS"), cur["STATS"]
    np.testing.assert_array_equal(ps[:, 0], st[:, 0])
    np.testing.assert_array_equal(ps[:, 1], st[:, 0])
    np.testing.assert_array_equal(ps[:, 2:], st[:, 2:])
    # the sensor prev rows are maintained by the step (updateSensorOutputIdx), not by shift
    np.testing.assert_array_equal(o.tensor("PREV_SENSOR_SEMANTIC"), sem)
    np.testing.assert_array_equal(o.tensor("PREV_SENSOR_DEPTH"), dep)
    o.shift_observations()   # idempotent (training_loop.py:135 calls it once per species)
    np.testing.assert_array_equal(o.tensor("PREV_POSITION"), cur["POSITION"])


def test_prev_sensor_rows_follow_their_agent():
    """updateSensorOutputIdx (src/sim/sim.cpp:771-784): after a step, the prev sensor row of a surviving agent
    is the row it had before the step, wherever the re-sort moved it."""
    o = OracleSim(6, 11, 32)
    _roll(o, 10, 6)
    rows0 = np.array(o.tensor("SENSOR_INDEX")[:, 0])
    wc0 = np.array(o.tensor("DEBUG_WORLD_COUNT")[:, 0])
    dep0, sem0 = np.array(o.tensor("SENSOR_DEPTH")), np.array(o.tensor("SENSOR_SEMANTIC"))
    o.tensor("ACTION")[:] = 0            # nobody moves, shoots or breeds: every agent survives in place
    pose0 = np.array(o.tensor("DEBUG_POSE"))
    o.step()
    assert np.array_equal(o.tensor("DEBUG_WORLD_COUNT")[:, 0] >= wc0, np.ones(6, bool))
    rows1 = o.tensor("SENSOR_INDEX")[:, 0]
    off0 = np.concatenate([[0], np.cumsum(wc0)[:-1]])
    off1 = o.tensor("DEBUG_WORLD_OFFSET")[:, 0]
    for w in range(6):
        for s in range(wc0[w]):
            np.testing.assert_array_equal(o.tensor("DEBUG_POSE")[off1[w] + s], pose0[off0[w] + s])
            r0, r1 = rows0[off0[w] + s], rows1[off1[w] + s]
            np.testing.assert_array_equal(o.tensor("PREV_SENSOR_DEPTH")[r1], dep0[r0])
            np.testing.assert_array_equal(o.tensor("PREV_SENSOR_SEMANTIC")[r1], sem0[r0])


def test_shard_invariance_of_the_oracle():
    """RNG streams are keyed on the GLOBAL world id (src/sim/sim.cpp:1238-1239): worlds [8, 16) simulated alone
    equal worlds [8, 16) of a 16-world run when both see the same per-agent actions."""
    full, part = OracleSim(16, 21, 32), OracleSim(8, 21, 32, world_id_offset=8)
    for t in range(40):
        for sim in (full, part):
            pos = sim.tensor("POSITION")
            idx = ((pos[:, 0] * 7.0 + pos[:, 1] * 3.0).astype(np.int64) + t) % 6
            a = np.zeros((sim.num_agents(), 6), dtype=np.int32)
            a[np.arange(len(idx)), idx] = 1
            sim.tensor("ACTION")[:] = a
            sim.step()
    off = full.tensor("DEBUG_WORLD_OFFSET")[8, 0]
    np.testing.assert_array_equal(full.tensor("DEBUG_WORLD_COUNT")[8:], part.tensor("DEBUG_WORLD_COUNT"))
    np.testing.assert_array_equal(full.tensor("DEBUG_POSE")[off:], part.tensor("DEBUG_POSE"))
    np.testing.assert_array_equal(full.tensor("DEBUG_AGENT_HEALTH")[off:], part.tensor("DEBUG_AGENT_HEALTH"))
    np.testing.assert_array_equal(full.tensor("DEBUG_AGENT_TARGET")[off:], part.tensor("DEBUG_AGENT_TARGET"))
    np.testing.assert_array_equal(full.tensor("SPECIES_COUNT")[8:], part.tensor("SPECIES_COUNT"))


def test_thread_count_does_not_change_results():
    lib = oracle_lib(False)
    digs = []
    for threads in (1, 3):
        lib.mbo_set_threads(threads)
        _, _, d = run_rollout(lambda w, s, i, c, o: OracleAdapter(w, s, i, cap=c, world_id_offset=o), ROLLOUTS["w5_init13"])
        digs.append(d)
    lib.mbo_set_threads(os.cpu_count() or 1)
    np.testing.assert_array_equal(digs[0], digs[1])


def test_population_bookkeeping():
    """count(t+1) = count(t) - deaths + births + respawns; species counts export = tracker counts before respawn."""
    o = OracleSim(10, 13, 32)
    prev_stats = o.episode_stats()
    prev_n = o.num_agents()
    rng = np.random.default_rng(0)
    for _ in range(60):
        o.tensor("ACTION")[:] = random_actions(rng, o.num_agents(), "bits")
        o.step()
        st = o.episode_stats()
        d = st - prev_stats
        assert o.num_agents() == prev_n - d[11] + d[10] + d[12]
        sc = o.tensor("SPECIES_COUNT")
        assert sc.sum() == o.num_agents() - d[12]
        assert st[2:6].tolist() == sc.sum(0).tolist()
        # respawn tops every species up to initAgents / 4 (src/sim/sim.cpp:825-836)
        species = np.array(o.tensor("DEBUG_AGENT_SPECIES")[:, 0])
        world = np.repeat(np.arange(10), o.tensor("DEBUG_WORLD_COUNT")[:, 0])
        cnt = np.zeros((10, 4), dtype=np.int64)
        np.add.at(cnt, (world, species - 1), 1)
        assert cnt.min() >= 8
        prev_stats, prev_n = st, o.num_agents()
