"""A `learn/`-compatible training step on top of SimManager.

The reference's loop (learn/training_loop.py:29-137) drives every species' network as

    logits, values, species_memory = model.forward(observations, prev_memory)      # :57-59
    new_memory = model.generate_memory(observations, species_memory)               # :60
    ...
    actor_loss, critic_loss = model.compute_loss(log_probs, rewards, prev_V, new_V)   # :95, learn/models.py:181-184

with `prev_memory` the simulator's 16-float hidden-state rows, but no committed network takes a memory argument
or has `generate_memory`: the loop does not run against learn/models.py. `MemoryA2C` is the
missing variant -- same feature MLP / recurrent cell / actor / critic shapes as learn/models.py:14-93, with the
16-float memory projected into the cell's initial state and the cell's output projected back to 16 floats --
and `train_epoch` is the loop body with the SimManager calls kept as the reference makes them (step, the tensor
accessors, construct_obs on cur and prev, shift_observations inside the species loop, action and memory write-back).

Positions taken (DEVIATIONS.md D5): the log-probability of the taken action is `log_softmax(logits)[action]` (the
reference indexes the raw logits, training_loop.py:94); sampling is torch's Categorical as in the reference, through
an optional torch.Generator so that a run replays. Training itself is torch autograd + a torch optimiser: plumbing
above the simulator, not a kernel path.
"""
import torch
from torch import nn

__all__ = ["MemoryA2C", "construct_obs", "train_epoch"]

OBS_DIM, HIDDEN_DIM, ACTION_DIM, MEMORY_DIM = 69, 128, 6, 16     # learn/training_loop.py:228-232, src/sim/types.hpp:37-39


class MemoryA2C(nn.Module):
    """forward(observations (n, 69), memory (n, 16)) -> (logits (n, 6), values (n, 1), species_memory (n, 128))."""

    def __init__(self, obs_dim=OBS_DIM, hidden_dim=HIDDEN_DIM, action_dim=ACTION_DIM, memory_dim=MEMORY_DIM,
                 rtype="GRU", activations=("Tanh", "ReLU")):
        super().__init__()
        if rtype not in ("GRU", "LSTM", "RNN"):
            raise ValueError("rtype must be GRU, LSTM or RNN")            # learn/models.py:71-78
        layers, width = [], obs_dim
        for name in activations:                                        # learn/models.py:56-69
            layers += [nn.Linear(width, hidden_dim), getattr(nn, name)()]
            width = hidden_dim
        self.feature = nn.Sequential(*layers)
        self.rtype = rtype
        self.cell = {"GRU": nn.GRUCell, "LSTM": nn.LSTMCell, "RNN": nn.RNNCell}[rtype](hidden_dim, hidden_dim)
        self.memory_in = nn.Linear(memory_dim, hidden_dim)
        self.memory_out = nn.Linear(hidden_dim, memory_dim)
        self.actor = nn.Sequential(nn.Linear(hidden_dim, hidden_dim), nn.ReLU(), nn.Linear(hidden_dim, action_dim))
        self.critic = nn.Sequential(nn.Linear(hidden_dim, hidden_dim), nn.ReLU(), nn.Linear(hidden_dim, 1))

    @property
    def device(self):                                                   # the loop reads model.device (:90)
        return next(self.parameters()).device

    def forward(self, observations, memory):
        x = self.feature(observations)
        h0 = torch.tanh(self.memory_in(memory))
        if self.rtype == "LSTM":
            h, _ = self.cell(x, (h0, torch.zeros_like(h0)))
        else:
            h = self.cell(x, h0)
        return self.actor(h), self.critic(h), h

    def generate_memory(self, observations, species_memory):
        """The 16 floats written back to the simulator's hidden-state rows; no gradient flows through the simulator."""
        return torch.tanh(self.memory_out(species_memory)).detach()

    @staticmethod
    def compute_loss(action_log_probs, reward, prev_V, new_V, critic_loss=None, gamma=1.0):
        """learn/models.py:181-184: one-step advantage actor loss (summed) and a smooth-L1 critic loss."""
        assert len(action_log_probs) == len(reward) == len(prev_V) == len(new_V)
        critic_loss = critic_loss or nn.SmoothL1Loss()
        advantage = reward + gamma * new_V.detach().squeeze() - prev_V.detach().squeeze()
        return -(torch.sum(action_log_probs * advantage)), critic_loss(reward, prev_V.squeeze())


def construct_obs(sim_mgr, start, end, prev=False):
    """learn/util.py:23-29: (end - start, 69) float32 = depth 32 | health 1 | position 2 | semantic 32 | surrounding 2.
    One kernel (`pack_observations`) on the CUDA SimManager, torch.cat on anything else with the tensor accessors."""
    start, end = int(start), int(end)
    if hasattr(sim_mgr, "pack_observations"):
        return sim_mgr.pack_observations(start, end, prev)
    parts = [sim_mgr.depth_tensor(prev), sim_mgr.health_tensor(prev), sim_mgr.position_tensor(prev),
             sim_mgr.semantic_tensor(prev), sim_mgr.surrounding_tensor(prev)]
    return torch.cat([p.to_torch()[start:end].float() for p in parts], dim=1)


def train_epoch(sim_mgr, species_nets, species_optims, generator=None):
    """One pass of learn/training_loop.py:36-136 (without logging and checkpointing): step the simulator, then per
    species act on the current observations, take one A2C gradient step on the previous transition, shift the
    observations and write the new actions and memories back. Returns [(actor_loss, critic_loss, agents)] * 4."""
    sim_mgr.step()
    if hasattr(sim_mgr, "species_blocks"):                               # exact row ranges
        starts, ends = zip(*sim_mgr.species_blocks())
    else:                                                                # the reference's way (:43-45): counts taken before the
        counts = sim_mgr.species_count_tensor().to_torch()              # step's respawns, so short by the respawned agents
        ends = counts.sum(dim=0).cumsum(dim=0).tolist()
        starts = [0] + ends[:-1]
    action_tensor = sim_mgr.action_tensor(False).to_torch()
    memory_tensor = sim_mgr.hidden_state_tensor(False).to_torch()
    all_rewards = sim_mgr.reward_tensor(False).to_torch().clone()
    out = []
    for k, (sp_start, sp_end) in enumerate(zip(starts, ends)):
        sp_start, sp_end = int(sp_start), int(sp_end)
        model, optimizer = species_nets[k], species_optims[k]
        if sp_end == sp_start:
            sim_mgr.shift_observations()
            out.append((0.0, 0.0, 0))
            continue
        dev = model.device
        observations = construct_obs(sim_mgr, sp_start, sp_end, prev=False).to(dev)
        prev_memory = memory_tensor[sp_start:sp_end, :].to(dev)
        logits, new_values, species_memory = model.forward(observations, prev_memory)
        new_memory = model.generate_memory(observations, species_memory)
        actions = torch.multinomial(torch.softmax(logits.detach().float(), dim=1), 1, generator=generator).squeeze(1)
        one_hot = torch.zeros(actions.size(0), ACTION_DIM, device=actions.device)
        one_hot.scatter_(1, actions.unsqueeze(1), 1)

        rewards = all_rewards[sp_start:sp_end, :].to(dev)
        prev_observations = construct_obs(sim_mgr, sp_start, sp_end, prev=True).to(dev)
        og_memory = sim_mgr.hidden_state_tensor(True).to_torch()[sp_start:sp_end, :].to(dev)
        prev_logits, prev_values, _ = model.forward(prev_observations, og_memory)
        prev_actions = action_tensor[sp_start:sp_end, :].argmax(dim=1).to(dev)      # the actions that led here
        prev_log_probs = torch.log_softmax(prev_logits, dim=1)[torch.arange(prev_actions.shape[0], device=dev), prev_actions]
        actor_loss, critic_loss = model.compute_loss(prev_log_probs, rewards.flatten(), prev_values.flatten(),
                                                     new_values.flatten())
        optimizer.zero_grad()
        (actor_loss + critic_loss).backward()
        optimizer.step()

        sim_mgr.shift_observations()                                    # inside the species loop, as the reference (:135)
        action_tensor[sp_start:sp_end, :] = one_hot.to(action_tensor.device).to(action_tensor.dtype)
        memory_tensor[sp_start:sp_end, :] = new_memory.to(memory_tensor.device)
        out.append((float(actor_loss.detach()), float(critic_loss.detach()), sp_end - sp_start))
    return out
