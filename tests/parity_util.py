"""Shared driver of the parity tests: advances the CPU oracle and the CUDA simulator with the
same seeds, actions and hidden states and compares every export after every call.

Tolerances: integer / byte exports must be bit-exact. Float exports (position, surrounding,
reward, health, hidden) are ALSO required to be bit-exact -- the kernels are built without FMA
contraction and use the shared deterministic sin/cos/atan2 of include/mbots_detmath.h, so
FLOAT_TOL = 0.0 is the stated tolerance; it is kept as a named constant so a platform that
cannot meet it can relax it in one place.
"""
import numpy as np

from oracle_binding import OracleSim, random_actions

FLOAT_TOL = 0.0

ROW_EXPORTS = [
    "ACTION", "PREV_ACTION", "HIDDEN_STATE", "PREV_HIDDEN_STATE", "REWARD", "PREV_REWARD",
    "POSITION", "PREV_POSITION", "HEALTH", "PREV_HEALTH", "SURROUNDING", "PREV_SURROUNDING",
    "SENSOR_SEMANTIC", "SENSOR_DEPTH", "PREV_SENSOR_SEMANTIC", "PREV_SENSOR_DEPTH", "STATS", "PREV_STATS",
    "SENSOR_INDEX", "DONE",
]
WORLD_EXPORTS = ["SPECIES_COUNT", "DEBUG_WORLD_COUNT", "DEBUG_WORLD_OFFSET", "DEBUG_RNG_COUNT",
                 "DEBUG_NUM_FOOD", "DEBUG_ROW_BASE"]
AGENT_DEBUG = ["DEBUG_POSE", "DEBUG_AGENT_HEALTH", "DEBUG_AGENT_SPECIES", "DEBUG_AGENT_TARGET"]


def gpu_np(sim, name):
    return sim.export(name).to_torch().cpu().numpy()


def food_canonical(raw):
    """(W*48, 8) uint8 -> the cells of packages that hold food (cells of empty packages are stale)."""
    raw = raw.copy()
    mask = raw[:, 5]
    for p in range(5):
        raw[:, p] = np.where((mask >> p) & 1, raw[:, p], 0)
    raw[:, 6:] = 0
    return raw


def compare(o, g, where, sensors=True):
    A = o.num_agents()
    assert g.num_agents() == A, f"{where}: agent count {g.num_agents()} != oracle {A}"
    for name in WORLD_EXPORTS:
        a, b = o.tensor(name), gpu_np(g, name)
        assert a.shape == b.shape, f"{where}: {name} shape {b.shape} != {a.shape}"
        np.testing.assert_array_equal(b, a, err_msg=f"{where}: {name}")
    np.testing.assert_array_equal(food_canonical(gpu_np(g, "DEBUG_FOOD")), food_canonical(o.tensor("DEBUG_FOOD")),
                                  err_msg=f"{where}: food")
    for name in AGENT_DEBUG + ROW_EXPORTS:
        a, b = o.tensor(name), gpu_np(g, name)
        assert a.shape == b.shape, f"{where}: {name} shape {b.shape} != {a.shape}"
        assert a.dtype == b.dtype, f"{where}: {name} dtype {b.dtype} != {a.dtype}"
        if a.dtype == np.float32 and FLOAT_TOL > 0:
            np.testing.assert_allclose(b, a, rtol=FLOAT_TOL, atol=FLOAT_TOL, err_msg=f"{where}: {name}")
        else:
            np.testing.assert_array_equal(b, a, err_msg=f"{where}: {name}")
    np.testing.assert_array_equal(np.array(g.episode_stats(), dtype=np.int64), o.episode_stats(),
                                  err_msg=f"{where}: episode stats")


def drive(o, g, steps, seed, mode="onehot", shift_every=1, check_every=1, write_hidden=True):
    """Advance both simulators `steps` times; returns the per-step agent counts."""
    import torch
    rng = np.random.default_rng(seed)
    counts = []
    compare(o, g, "init")
    for t in range(steps):
        A = o.num_agents()
        acts = random_actions(rng, A, mode)
        o.tensor("ACTION")[:] = acts
        if A:
            g.action_tensor(False).to_torch().copy_(torch.from_numpy(acts))
        if write_hidden and A:
            hid = rng.standard_normal((A, 16)).astype(np.float32)
            o.tensor("HIDDEN_STATE")[:] = hid
            g.hidden_state_tensor(False).to_torch().copy_(torch.from_numpy(hid))
        o.step()
        g.step()
        counts.append(o.num_agents())
        if (t + 1) % check_every == 0 or t == steps - 1:
            compare(o, g, f"step {t}")
        if shift_every and (t + 1) % shift_every == 0:
            o.shift_observations()
            g.shift_observations()
            if (t + 1) % check_every == 0 or t == steps - 1:
                compare(o, g, f"shift after step {t}")
    return counts


def make_pair(num_worlds, seed, init_agents, cap=0, world_id_offset=0, gpu_id=0):
    from madrona_bots import SimManager
    o = OracleSim(num_worlds, seed, init_agents, cap=cap, world_id_offset=world_id_offset)
    g = SimManager(gpu_id, num_worlds, seed, init_agents, max_agents_per_world=cap, world_id_offset=world_id_offset)
    return o, g
