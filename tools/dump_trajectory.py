#!/usr/bin/env python
"""Random-action rollout with a headless trajectory dump (stands in for the reference's viewer):
    python tools/dump_trajectory.py --worlds 64 --steps 200 --record 0,1,2 --out rollout.npz"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "madrona-bots_b200"))


def main():
    import torch
    from madrona_bots import SimManager
    from madrona_bots.trajectory import TrajectoryRecorder
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, default=64)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--record", default="0", help="comma-separated world ids")
    ap.add_argument("--out", default="rollout.npz")
    a = ap.parse_args()
    sim = SimManager(0, a.worlds, a.seed, 32)
    rec = TrajectoryRecorder(sim, worlds=[int(x) for x in a.record.split(",")])
    gen = torch.Generator(device="cuda").manual_seed(1234)
    for _ in range(a.steps):
        act = sim.action_tensor(False).to_torch()
        act.zero_()
        act.scatter_(1, torch.randint(0, 6, (act.shape[0], 1), generator=gen, device="cuda"), 1)
        sim.step()
        rec.record()
        sim.shift_observations()
    rec.save(a.out)
    print(f"wrote {a.out}: {a.steps} steps, worlds {a.record}")


if __name__ == "__main__":
    main()
