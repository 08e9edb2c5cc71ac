"""Soak of the no-host-wait acting path: two handles with the shipped universe_violence networks in the loop, one driven by the
blocking loop (step, act through species_blocks(), shift), one by act_async (mbots_policy_act_all) + rollout_step_async with the
host `depth` steps ahead; every `check` steps all exports must be identical. One JSON line.
    python tools/soak_policy.py WORLDS STEPS [CHECK] [DEPTH]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "madrona-bots_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


def main():
    import numpy as np
    import torch
    from madrona_bots import SimManager
    from madrona_bots.policy import PolicyRunner, load_universe_fixture
    from parity_util import ROW_EXPORTS, WORLD_EXPORTS, AGENT_DEBUG, gpu_np
    W, steps = int(sys.argv[1]), int(sys.argv[2])
    check = int(sys.argv[3]) if len(sys.argv) > 3 else 50
    depth = int(sys.argv[4]) if len(sys.argv) > 4 else 4
    pols = load_universe_fixture(os.path.join(ROOT, "tests", "golden", "policy_universe_violence.npz"), "cuda:0")
    a, b = SimManager(0, W, 7, 32), SimManager(0, W, 7, 32)
    ra, rb = PolicyRunner(a, pols, seed=5, fused="sw128"), PolicyRunner(b, pols, seed=5, fused="sw128")
    for s in (a, b):
        s.step(); s.shift_observations()
    agent_steps, checks, inflight = 0, 0, []
    for t in range(1, steps + 1):
        ra.act(); a.step(); a.shift_observations()
        rb.act_async(); b.rollout_step_async(None, shift=True)
        ev = torch.cuda.Event(); ev.record(); inflight.append(ev)
        if len(inflight) > depth:
            inflight.pop(0).synchronize()
        agent_steps += a.num_agents()
        if t % check == 0 or t == steps:
            b.sync()
            assert a.num_agents() == b.num_agents(), f"step {t}: {a.num_agents()} vs {b.num_agents()} agents"
            for name in WORLD_EXPORTS + AGENT_DEBUG + ROW_EXPORTS + ["DEBUG_FOOD"]:
                np.testing.assert_array_equal(gpu_np(a, name), gpu_np(b, name), err_msg=f"step {t}: {name}")
            checks += 1
    st = b.episode_stats()
    print(json.dumps({"worlds": W, "steps": steps, "agent_steps": agent_steps, "checks": checks, "host_steps_ahead": depth,
                      "births": st[10], "deaths": st[11], "respawns": st[12], "food_eaten": st[14], "shots_hit": st[15],
                      "agents_at_end": b.num_agents(), "result": "identical on every export at every check"}))


if __name__ == "__main__":
    main()
