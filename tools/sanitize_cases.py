"""Workloads run under compute-sanitizer (tools/gpu_sanitize.sh). Each case drives the CUDA path only (the
oracle comparison is the parity tests' job); sizes are small because the sanitizer slows kernels 10-100x."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "madrona-bots_b200"), os.path.join(ROOT, "tests"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np
import torch


def rollout(worlds, init, steps, cap=0, mode="bits", seed=3):
    from madrona_bots import SimManager
    from oracle_binding import random_actions
    sim = SimManager(0, worlds, seed, init, max_agents_per_world=cap)
    rng = np.random.default_rng(seed)
    for _ in range(steps):
        A = sim.num_agents()
        if A:
            sim.action_tensor(False).to_torch().copy_(torch.from_numpy(random_actions(rng, A, mode)))
        sim.step()
        sim.shift_observations()
    wc = sim.export("DEBUG_WORLD_COUNT").to_torch()
    return sim, int(wc.max()), sim.episode_stats()


def main():
    case = sys.argv[1]
    if case == "smoke":
        import __graft_entry__ as g
        g.smoke()
    elif case == "w1024":
        sim, mx, st = rollout(1024, 32, 10, mode="onehot")
        print("max world", mx, "stats", st[:2], st[10:14])
    elif case == "full":
        # worlds above 64 agents: the full export variant (64-bit keys, 8-observer tiles) and the step kernel's
        # third pass; 44 initial agents breed across 64 within a few steps
        sim, mx, st = rollout(24, 60, 12, cap=96)
        assert mx > 64, mx
        print("max world", mx, "stats", st[:2], st[10:14])
    elif case == "cap32":
        # capacity 32 (no second slot batch): 4096 worlds x 32 slots x 16 B = exactly 2 MiB of pose
        sim, mx, st = rollout(4096, 32, 4, cap=32)
        print("max world", mx, "overflow", st[13])
    elif case == "policy":
        from madrona_bots import SimManager
        from madrona_bots.policy import FusedPolicy, SpeciesPolicy
        from test_policy import _make
        sim = SimManager(0, 64, 1, 32)
        sim.step()
        blocks = sim.species_blocks()
        nets = (("GRU", ["Tanh", "ReLU"]), ("LSTM", ["LeakyReLU", "ReLU"]), ("RNN", ["LogSigmoid", "Tanh", "ELU"]),
                ("RNN", ["LogSigmoid", "Tanh", "LeakyReLU"]))
        for layout in ("sw128", "rowpad"):
            for k, (rtype, acts) in enumerate(nets):
                torch.manual_seed(k)
                net, config = _make(rtype, acts)
                fused = FusedPolicy(SpeciesPolicy(config, net.state_dict(), "cuda:0"), layout)
                a, b = blocks[k]
                sim.policy_act(a, b, fused, seed=5, draw=1, want_index=True, want_logits=True)
        torch.cuda.synchronize()
    else:
        raise SystemExit(f"unknown case {case}")
    torch.cuda.synchronize()
    print("case ok:", case)


if __name__ == "__main__":
    main()
