#!/usr/bin/env python
"""Generates tests/golden/refsim_rollouts.npz -- golden vectors produced by the REFERENCE'S OWN SOURCE TEXT.

Run where /root/reference is mounted:   python tests/golden/make_refsim_golden.py

oracle/_ref/librefsim.so (recipe: `make -C oracle refsim`) compiles src/sim/sim.cpp, external/madrona/src/render/
ecs_system.cpp and the raycast functions of external/madrona/src/mw/device/bvh_raycast.cpp from where they lie
and runs ONE world through the reference's own task graphs under the host mock of oracle/refmock/ (rules U5, U6,
U8, U19, D2, D3 of DEVIATIONS.md applied, each switchable). For every case below the fixture stores, per step,
the agent count and a CRC32 of every export the reference world can be asked for (refsim_binding.REF_EXPORTS),
in the export formats of include/mbots.h, plus the complete final state of the first case.

The GPU box has no /root/reference: there the CPU oracle (tests/test_reference_bodies.py) and the CUDA path
(tests/test_parity_gpu.py::test_cuda_path_reproduces_reference_text_rollouts) are checked against this file.
"""
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle_binding import random_actions  # noqa: E402
from refsim_binding import REF_EXPORTS, RefSimAdapter  # noqa: E402

# name -> (seed, world id, init agents, steps, action mode, action seed)
CASES = {
    "s0_w0": (0, 0, 32, 150, "onehot", 1),
    "s7_w3_bits": (7, 3, 32, 150, "bits", 2),
    "s21_w4099": (21, 4099, 32, 100, "onehot", 3),
    "s9_w1_init13": (9, 1, 13, 100, "bits", 4),
    "s5_w2_init48": (5, 2, 48, 60, "bits", 5),
}


def food_canonical(raw):
    raw = np.array(raw, copy=True)
    mask = raw[:, 5]
    for p in range(5):
        raw[:, p] = np.where((mask >> p) & 1, raw[:, p], 0)
    raw[:, 6:] = 0
    return raw


def digest(sim):
    out = []
    for n in REF_EXPORTS:
        t = sim.tensor(n)
        if n == "DEBUG_FOOD":
            t = food_canonical(t)
        out.append(zlib.crc32(np.ascontiguousarray(t).tobytes()))
    return np.array(out, dtype=np.uint32)


def run_case(make_sim, cfg, on_step=None):
    """make_sim(seed, world_id, init) -> object with the host-array surface (tensor / write / step / shift_observations /
    num_agents). Returns (sim, counts, digests): digests[t] is taken after step t (and its shift, two steps in three)."""
    seed, wid, init, steps, mode, aseed = cfg
    sim = make_sim(seed, wid, init)
    rng = np.random.default_rng(aseed)
    counts, digs = [sim.num_agents()], []
    for t in range(steps):
        A = sim.num_agents()
        sim.write("ACTION", random_actions(rng, A, mode))
        sim.write("HIDDEN_STATE", rng.standard_normal((A, 16)).astype(np.float32))
        sim.step()
        if t % 3 != 2:
            sim.shift_observations()
        counts.append(sim.num_agents())
        digs.append(digest(sim))
        if on_step:
            on_step(t, sim)
    return sim, np.array(counts, dtype=np.int64), np.stack(digs)


def main():
    out = {}
    for name, cfg in CASES.items():
        sim, counts, digs = run_case(lambda s, w, i: RefSimAdapter(s, w, i), cfg)
        out[name + ".counts"] = counts
        out[name + ".digests"] = digs
        if name == "s0_w0":
            for n in REF_EXPORTS:
                out["s0_w0.final." + n] = np.array(sim.tensor(n))
        print(name, "agents", counts[0], "->", counts[-1], "min", counts.min(), "max", counts.max())
    np.savez_compressed(os.path.join(HERE, "refsim_rollouts.npz"), **out)
    print("wrote refsim_rollouts.npz")


if __name__ == "__main__":
    main()
