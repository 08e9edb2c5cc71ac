#!/usr/bin/env python
"""Long parity soak: CUDA path vs CPU oracle over thousands of steps (every export compared every N steps).
    python tools/soak.py --worlds 256 --steps 3000 --mode bits --check-every 250"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "madrona-bots_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


def drive_fused(o, g, steps, seed, mode, check_every, depth=4):
    """The oracle steps on the CPU while up to `depth` fused graph launches are in flight on the GPU."""
    import numpy as np
    import torch
    from oracle_binding import random_actions
    from parity_util import compare
    rng = np.random.default_rng(seed)
    counts, pend = [], []
    for t in range(steps):
        A = o.num_agents()
        acts = random_actions(rng, A, mode)
        o.tensor("ACTION")[:] = acts
        o.step(); o.shift_observations()
        dev = torch.from_numpy(acts).cuda()
        g.rollout_step_async(dev, shift=True)
        pend.append(dev)
        counts.append(o.num_agents())
        if len(pend) >= depth or (t + 1) % check_every == 0 or t == steps - 1:
            g.sync()
            torch.cuda.synchronize()
            pend.clear()
        if (t + 1) % check_every == 0 or t == steps - 1:
            compare(o, g, f"fused step {t}")
    return counts


def main():
    from parity_util import drive, make_pair
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, default=256)
    ap.add_argument("--steps", type=int, default=3000)
    ap.add_argument("--mode", default="bits", choices=["bits", "onehot"])
    ap.add_argument("--check-every", type=int, default=250)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--cap", type=int, default=0)
    ap.add_argument("--fused", action="store_true",
                    help="drive the CUDA side through mbots_rollout_step_async (one graph launch per step: actions from a "
                         "device matrix, fused shift), several steps in flight between checks")
    a = ap.parse_args()
    o, g = make_pair(a.worlds, a.seed, 32, cap=a.cap)
    t0 = time.time()
    if a.fused:
        counts = drive_fused(o, g, a.steps, a.seed + 1, a.mode, a.check_every)
    else:
        counts = drive(o, g, steps=a.steps, seed=a.seed + 1, mode=a.mode, check_every=a.check_every)
    st = [int(x) for x in o.episode_stats()]
    print(json.dumps({"worlds": a.worlds, "steps": a.steps, "mode": a.mode, "check_every": a.check_every,
                      "path": "fused graph step, run-ahead 4" if a.fused else "blocking step + shift",
                      "result": "every export bit-identical at every check", "agents_min": min(counts),
                      "agents_max": max(counts), "agent_steps": st[0], "births": st[10], "deaths": st[11],
                      "respawns": st[12], "overflow": st[13], "eaten": st[14], "shots_hit": st[15],
                      "seconds": round(time.time() - t0, 1)}))


if __name__ == "__main__":
    main()
