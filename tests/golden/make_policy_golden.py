#!/usr/bin/env python
"""Generates tests/golden/policy_universe_violence.npz -- the reference's OWN Python policy on the shipped weights.

Run where /root/reference is mounted:   python tests/golden/make_policy_golden.py

Imports /root/reference/learn/models.py (ActorCritic, learn/models.py:95-165), loads
checkpoints/universe_violence/species_{1..4}/latest_model_epoch_788.pt (weights_only=True) and feeds
512 fixed observation rows ONE AGENT AT A TIME: `forward` on a (1, 69) matrix is the unbatched one-element sequence of
the recurrent layer from a zero state -- exactly the "one recurrent step per agent from a zero state" that the acting
path implements (DEVIATIONS.md D4). Stored: the observation rows (already as the five exports that make them up), the
fp32 logits and values of the reference's modules, and the four networks' weights + configs (fp32, 1.9 MB) so that the
GPU box -- which has no /root/reference -- can rebuild the FusedPolicy blobs and run `tools/policy_rollout.py --universe`.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
N = 512


def observations(seed=20261105):
    """Rows shaped like the simulator's exports: depth uint8 x32 (255 = miss, pixel 0 = row stamp), health float
    (numeric, rule U4), position in the 128 x 96 world, semantic int8 x32 (-1 = miss, 1..4), surroundings (bilinear
    chunk statistics)."""
    rng = np.random.default_rng(seed)
    hit = rng.random((N, 32)) < 0.45
    depth = np.where(hit, rng.integers(1, 160, (N, 32)), 255).astype(np.uint8)
    depth[:, 0] = (np.arange(N) & 0xFF).astype(np.uint8)
    semantic = np.where(hit, rng.integers(1, 5, (N, 32)), -1).astype(np.int8)
    health = rng.choice([10, 30, 50, 60, 80, 100, 120, 140], size=(N, 1)).astype(np.float32)
    position = np.stack([rng.random(N) * 127.0, rng.random(N) * 95.0], axis=1).astype(np.float32)
    surrounding = np.stack([rng.random(N) * 3.0, rng.random(N) * 6.0], axis=1).astype(np.float32)
    return depth, health, position, semantic, surrounding


def construct_obs(depth, health, position, semantic, surrounding):
    """learn/util.py:23-29: torch.cat((depth, health, position, semantic, surrounding), dim=1) as float32."""
    return np.concatenate([depth.astype(np.float32), health, position, semantic.astype(np.float32), surrounding], axis=1)


def main():
    sys.path.insert(0, os.path.join(REF, "learn"))
    from models import ActorCritic        # the reference's module, imported from where it lies
    parts = observations()
    obs = torch.from_numpy(construct_obs(*parts))
    out = {"depth": parts[0], "health": parts[1], "position": parts[2], "semantic": parts[3], "surrounding": parts[4]}
    for k in range(1, 5):
        path = os.path.join(REF, "checkpoints", "universe_violence", f"species_{k}", "latest_model_epoch_788.pt")
        ck = torch.load(path, map_location="cpu", weights_only=True)
        cfg = ck["model_config"]
        model = ActorCritic(69, 6, None, "cpu", config=cfg)
        model.load_state_dict(ck["model_state_dict"])
        model.eval()
        logits, values = [], []
        with torch.no_grad():
            for i in range(N):
                lg, v = model.forward(obs[i:i + 1])           # one agent = a one-element sequence from a zero state
                logits.append(lg[0].numpy())
                values.append(v[0].numpy())
            seq_logits, _ = model.forward(obs)                 # what the reference's loop literally does: ONE sequence
        out[f"sp{k}.logits"] = np.stack(logits).astype(np.float32)
        out[f"sp{k}.values"] = np.stack(values).astype(np.float32)
        out[f"sp{k}.logits_as_one_sequence_row0"] = seq_logits[0].numpy().astype(np.float32)
        out[f"sp{k}.config_json"] = np.frombuffer(__import__("json").dumps(cfg).encode(), dtype=np.uint8)
        for name, t in ck["model_state_dict"].items():
            out[f"sp{k}.w.{name}"] = t.numpy().astype(np.float32)
        print(f"species {k}: {cfg['recurrent']['type']}, {sum(1 for l in cfg['layers'] if l['type'] == 'linear')} feature layers, "
              f"logit scale {np.abs(out[f'sp{k}.logits']).max():.3f}, row 0 of the one-sequence forward == per-agent row 0: "
              f"{np.allclose(out[f'sp{k}.logits'][0], out[f'sp{k}.logits_as_one_sequence_row0'], atol=1e-6)}")
    np.savez_compressed(os.path.join(HERE, "policy_universe_violence.npz"), **out)
    print("wrote policy_universe_violence.npz", os.path.getsize(os.path.join(HERE, "policy_universe_violence.npz")) // 1024, "KB")


if __name__ == "__main__":
    main()
