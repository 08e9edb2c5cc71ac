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


def main():
    from parity_util import drive, make_pair
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, default=256)
    ap.add_argument("--steps", type=int, default=3000)
    ap.add_argument("--mode", default="bits", choices=["bits", "onehot"])
    ap.add_argument("--check-every", type=int, default=250)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--cap", type=int, default=0)
    a = ap.parse_args()
    o, g = make_pair(a.worlds, a.seed, 32, cap=a.cap)
    t0 = time.time()
    counts = drive(o, g, steps=a.steps, seed=a.seed + 1, mode=a.mode, check_every=a.check_every)
    st = [int(x) for x in o.episode_stats()]
    print(json.dumps({"worlds": a.worlds, "steps": a.steps, "mode": a.mode, "check_every": a.check_every,
                      "result": "every export bit-identical at every check", "agents_min": min(counts),
                      "agents_max": max(counts), "agent_steps": st[0], "births": st[10], "deaths": st[11],
                      "respawns": st[12], "overflow": st[13], "eaten": st[14], "shots_hit": st[15],
                      "seconds": round(time.time() - t0, 1)}))


if __name__ == "__main__":
    main()
