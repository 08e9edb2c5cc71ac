#!/usr/bin/env python
"""Generates the committed golden fixtures of tests/golden/ -- run in the container where
/root/reference is mounted:

    python tests/golden/make_golden.py

  prims_ref.npz   outputs of the reference's OWN rand.hpp / math.hpp (through oracle/_ref, the oracle
                  compiled with -DMBO_REF_PRIMS against the headers where they lie under /root/reference)
                  on a fixed table of inputs: Threefry split_i, sampleI32, sampleUniform, and a 1,500-turn
                  quaternion walk with rotateVec / normalize.
  rollout_ref.npz trajectory digests of the whole step (oracle/_ref flavour): for a few small
                  configurations, per step the agent count and a CRC32 of every export, plus the complete
                  final state of the smallest one. The GPU box has no /root/reference; there the restated
                  oracle and the CUDA path are both checked against this file.

The reference holds no golden vectors for the step itself; what these fixtures pin
is (a) the restated primitives to the reference's code and (b) today's oracle semantics against silent
drift, NOT the step semantics to a run of the reference (it cannot run: SURVEY.md section 0).
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle_binding import OracleSim, oracle_lib, random_actions  # noqa: E402

DIGEST_EXPORTS = [
    "ACTION", "PREV_ACTION", "HIDDEN_STATE", "PREV_HIDDEN_STATE", "REWARD", "PREV_REWARD", "POSITION",
    "PREV_POSITION", "HEALTH", "PREV_HEALTH", "SURROUNDING", "PREV_SURROUNDING", "SENSOR_SEMANTIC",
    "SENSOR_DEPTH", "PREV_SENSOR_SEMANTIC", "PREV_SENSOR_DEPTH", "STATS", "PREV_STATS", "SENSOR_INDEX",
    "SPECIES_COUNT", "DEBUG_WORLD_COUNT", "DEBUG_RNG_COUNT", "DEBUG_NUM_FOOD", "DEBUG_POSE",
    "DEBUG_AGENT_HEALTH", "DEBUG_AGENT_SPECIES", "DEBUG_AGENT_TARGET", "DEBUG_ROW_BASE",
]

# name -> (worlds, seed, init agents, capacity, world id offset, steps, action mode, action seed)
ROLLOUTS = {
    "w1": (1, 0, 32, 0, 0, 200, "onehot", 1),
    "w24_bits": (24, 7, 32, 0, 0, 80, "bits", 2),
    "w16_offset": (16, 21, 32, 0, 4096, 60, "onehot", 3),
    "w8_cap32": (8, 3, 32, 32, 0, 80, "bits", 4),
    "w5_init13": (5, 9, 13, 0, 0, 60, "bits", 5),
}


def digest(sim):
    """CRC32 of every export of a simulator exposing .tensor(name) -> numpy."""
    return np.array([zlib.crc32(np.ascontiguousarray(sim.tensor(n)).tobytes()) for n in DIGEST_EXPORTS],
                    dtype=np.uint32)


def run_rollout(make_sim, cfg, on_step=None):
    """Drives any simulator with the oracle's host-array surface; returns (counts, digests)."""
    worlds, seed, init, cap, off, steps, mode, aseed = cfg
    sim = make_sim(worlds, seed, init, cap, off)
    rng = np.random.default_rng(aseed)
    counts, digs = [sim.num_agents()], [digest(sim)]
    for t in range(steps):
        A = sim.num_agents()
        sim.write("ACTION", random_actions(rng, A, mode))
        sim.write("HIDDEN_STATE", rng.standard_normal((A, 16)).astype(np.float32))
        sim.step()
        if t % 3 != 2:                       # leave some steps unshifted: prev columns must be carried, not rebuilt
            sim.shift_observations()
        counts.append(sim.num_agents())
        digs.append(digest(sim))
        if on_step:
            on_step(t, sim)
    return sim, np.array(counts, dtype=np.int64), np.stack(digs)


class OracleAdapter(OracleSim):
    def write(self, name, value):
        if len(value):
            self.tensor(name)[:] = value


def main():
    from test_oracle_prims import _prim_table
    ref = oracle_lib(True)
    if ref is None:
        sys.exit("oracle/_ref is unavailable: run where /root/reference is mounted")
    assert ref.mbo_flavour() == b"reference-headers"
    np.savez_compressed(os.path.join(HERE, "prims_ref.npz"), **_prim_table(ref))
    out = {}
    for name, cfg in ROLLOUTS.items():
        sim, counts, digs = run_rollout(lambda w, s, i, c, o: OracleAdapter(w, s, i, cap=c, world_id_offset=o, ref=True), cfg)
        out[name + ".counts"] = counts
        out[name + ".digests"] = digs
        out[name + ".stats"] = sim.episode_stats()
        if name == "w1":
            for n in DIGEST_EXPORTS:
                out["w1.final." + n] = np.array(sim.tensor(n))
    np.savez_compressed(os.path.join(HERE, "rollout_ref.npz"), **out)
    print("wrote", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
