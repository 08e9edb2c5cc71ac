"""Headless trajectory dump.

The reference inspects a running simulation through its Vulkan viewer and the "Raycast Visualizer"
window (src/gfx/gfx.cpp:157-318): agent positions coloured by species, and per agent the 32 depth /
semantic sensor pixels. No graphics stack exists here, so the same data is written to an `.npz`
instead -- per recorded step, for a chosen set of worlds, every live agent's pose, species, health,
reward, finder target and sensor rows -- for offline plotting. Device buffers only; no graphics API.
"""
import numpy as np

__all__ = ["TrajectoryRecorder"]


class TrajectoryRecorder:
    """Usage:
        rec = TrajectoryRecorder(sim, worlds=[0, 1, 2])
        for _ in range(T): ...write actions...; sim.step(); rec.record()
        rec.save("rollout.npz")

    Arrays in the file (N = agents recorded over all steps, concatenated step after step):
        step_offsets (T+1,) int64   agents of step t are rows [step_offsets[t], step_offsets[t+1])
        world (N,) int32, slot (N,) int32, row (N,) int32 (observation row = index into the exports)
        pose (N,4) float32 x, y, quat.w, quat.z;  species (N,) int8;  health (N,) int32
        target (N,) int32 finder hit as a slot of the same world, -1 none (bvh_raycast.cpp:272-293)
        reward (N,) float32;  action (N,6) int32 (what the agent was given for this step)
        depth (N,32) uint8, semantic (N,32) int8 (gfx.cpp:293,306 colours semantic+1)
        worlds (K,) int32 the recorded world ids; species_count (T,K,4) int32
    """

    def __init__(self, sim, worlds=None, sensors=True):
        import torch
        self.sim = sim
        self.sensors = bool(sensors)
        W = sim.num_worlds
        ids = list(range(min(W, 4))) if worlds is None else [int(w) for w in worlds]
        if any(w < 0 or w >= W for w in ids):
            raise ValueError("world id out of range")
        self._dev = torch.device("cuda", sim._gpu_id)
        self._worlds = torch.tensor(ids, dtype=torch.long, device=self._dev)
        self._steps = []

    def record(self):
        import torch
        sim = self.sim
        wl = self._worlds
        count = sim.export("DEBUG_WORLD_COUNT").to_torch()[:, 0].long()[wl]
        offset = sim.export("DEBUG_WORLD_OFFSET").to_torch()[:, 0].long()[wl]
        n = int(count.sum())
        # agent-order index of every live agent of the chosen worlds
        seg = torch.repeat_interleave(torch.arange(len(wl), device=self._dev), count)
        first = torch.cumsum(count, 0) - count
        slot = torch.arange(n, device=self._dev) - first[seg]
        idx = offset[seg] + slot
        rows = sim.sensor_index_tensor().to_torch()[:, 0].long()[idx]
        rec = {
            "world": wl[seg].to(torch.int32), "slot": slot.to(torch.int32), "row": rows.to(torch.int32),
            "pose": sim.export("DEBUG_POSE").to_torch()[idx].clone(),
            "species": sim.export("DEBUG_AGENT_SPECIES").to_torch()[idx, 0].to(torch.int8),
            "health": sim.export("DEBUG_AGENT_HEALTH").to_torch()[idx, 0].clone(),
            "target": sim.export("DEBUG_AGENT_TARGET").to_torch()[idx, 0].clone(),
            "reward": sim.reward_tensor(False).to_torch()[rows, 0],
            "action": sim.action_tensor(False).to_torch()[rows],
            "species_count": sim.species_count_tensor().to_torch()[wl],
        }
        if self.sensors:
            rec["depth"] = sim.depth_tensor(False).to_torch()[rows]
            rec["semantic"] = sim.semantic_tensor(False).to_torch()[rows]
        self._steps.append({k: v.cpu().numpy() for k, v in rec.items()})
        return n

    def arrays(self):
        steps = self._steps
        out = {"worlds": self._worlds.cpu().numpy().astype(np.int32),
               "step_offsets": np.concatenate([[0], np.cumsum([len(s["world"]) for s in steps])]).astype(np.int64)}
        if not steps:
            return out
        for k in steps[0]:
            if k == "species_count":
                out[k] = np.stack([s[k] for s in steps])
            else:
                out[k] = np.concatenate([s[k] for s in steps])
        return out

    def save(self, path):
        np.savez_compressed(path, **self.arrays())
