#!/usr/bin/env python
"""Simulator + species policies in the loop (BASELINE.json configs[4] shape, inference part): per step
sim.step(), PolicyRunner.act() (pack -> 4 species networks -> sample -> one-hot write), shift_observations().
Networks: the four architectures of checkpoints/universe_violence with random weights (the checkpoint itself
does not travel to the GPU box), or --universe DIR to load real ones.

    python tools/policy_rollout.py --worlds 8192 --steps 50
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "madrona-bots_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


def main():
    import torch
    from madrona_bots import SimManager
    from madrona_bots.policy import PolicyRunner, SpeciesPolicy, load_universe
    ap = argparse.ArgumentParser()
    ap.add_argument("--worlds", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--universe", default=None)
    ap.add_argument("--dtype", default="fp32", choices=["fp32", "tf32", "bf16"])
    ap.add_argument("--fused", nargs="?", const="sw128", default=None, choices=["sw128", "rowpad"],
                    help="one tensor-core kernel per species (mbots_policy_act): sw128 = tcgen05 kernel, rowpad = mma.sync kernel")
    a = ap.parse_args()
    if a.universe:
        pols = load_universe(a.universe, "cuda:0", dtype=a.dtype)
    else:
        from test_policy import _make
        torch.manual_seed(0)
        pols = []
        for rtype, acts in (("GRU", ["Tanh", "ReLU"]), ("LSTM", ["LeakyReLU", "ReLU"]),
                            ("RNN", ["LogSigmoid", "Tanh", "ReLU"]), ("RNN", ["LogSigmoid", "Tanh", "LeakyReLU"])):
            net, cfg = _make(rtype, acts)
            pols.append(SpeciesPolicy(cfg, net.state_dict(), "cuda:0", a.dtype))
    sim = SimManager(0, a.worlds, 0, 32)
    runner = PolicyRunner(sim, pols, seed=0, fused=a.fused)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    t_sim = t_pol = 0.0
    agent_steps = 0
    for t in range(a.warmup + a.steps):
        ev[0].record()
        sim.step()
        ev[1].record()
        runner.act()
        ev[2].record()
        sim.shift_observations()
        ev[3].record()
        torch.cuda.synchronize()
        if t >= a.warmup:
            t_sim += ev[0].elapsed_time(ev[1]) + ev[2].elapsed_time(ev[3])
            t_pol += ev[1].elapsed_time(ev[2])
            agent_steps += sim.num_agents()
    out = {"worlds": a.worlds, "steps": a.steps, "agents": sim.num_agents(),
           "sim_ms_per_step": t_sim / a.steps, "policy_ms_per_step": t_pol / a.steps,
           "agent_steps_per_s": agent_steps / ((t_sim + t_pol) * 1e-3),
           "policy": "universe_violence architectures, " + ("loaded" if a.universe else "random weights") + ", " + ("fused bf16 kernel (" + a.fused + ")" if a.fused else a.dtype)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
