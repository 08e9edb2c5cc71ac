"""Policy inference on the simulator's exports.

What the reference does per epoch and species (learn/training_loop.py:52-83,136, learn/util.py:23-29,
learn/models.py:95-165): concatenate five observation tensors into a (n, 69) matrix, run the species'
network -- feature MLP, one recurrent layer, actor head -- sample a Categorical action and write it back
one-hot into the action tensor. Here the two memory-bound ends are single CUDA kernels of libmbots.so
(`SimManager.pack_observations`, `SimManager.sample_actions`); the dense layers in between are plain
library GEMMs through torch.

Network definition = the checkpoint format of learn/ckpt.py:28-32 (`model_config` + `model_state_dict`
with the `a2c_nets.{feature,recurrent,actor,critic}` names of learn/models.py:6-12), so
checkpoints/universe_violence/species_*/latest_model_epoch_*.pt load as they are.

Position taken on the reference's caller drift: learn/models.py feeds the (n, 128)
feature matrix to nn.GRU/LSTM/RNN with no batch dimension and no state, which makes torch treat the n agents
of a species as ONE sequence; learn/training_loop.py instead expects a per-agent memory of 16 floats that no
committed network accepts. Neither is usable as a contract. This module runs the recurrent layer as one step
per agent from a zero state (agents are a batch), which is what "one step of a recurrent policy per agent per
simulator step" means and what the weights' shapes allow; the simulator's 16-float hidden-state tensor is
left untouched. Parity for this part is against torch's own modules (tests/test_policy.py), not the reference.
"""
import glob
import os

__all__ = ["SpeciesPolicy", "PolicyRunner", "load_universe"]

_ACTIVATIONS = ("Tanh", "ELU", "LogSigmoid", "LeakyReLU", "ReLU")     # learn/models.py:63-69


def _act(name, x):
    import torch.nn.functional as F
    import torch
    if name == "Tanh":
        return torch.tanh(x)
    if name == "ELU":
        return F.elu(x)
    if name == "LogSigmoid":
        return F.logsigmoid(x)
    if name == "LeakyReLU":
        return F.leaky_relu(x)
    if name == "ReLU":
        return F.relu(x)
    raise ValueError(f"unknown activation {name}")


class SpeciesPolicy:
    """One species' actor (and critic) from a reference checkpoint: inference only, fp32."""

    def __init__(self, config, state_dict, device):
        import torch
        self.config = config
        self.device = torch.device(device)
        sd = {k: v.detach().to(self.device, torch.float32).contiguous() for k, v in state_dict.items()}
        self.feature = self._stack(config["layers"], sd, "a2c_nets.feature")
        self.actor = self._stack(config["actor"], sd, "a2c_nets.actor")
        self.critic = self._stack(config["critic"], sd, "a2c_nets.critic")
        self.rtype = config["recurrent"]["type"]
        if self.rtype not in ("GRU", "LSTM", "RNN"):
            raise ValueError(f"unknown recurrent layer {self.rtype}")
        H = int(config["recurrent"]["hidden_dim"])
        self.hidden_dim = H
        w_ih = sd["a2c_nets.recurrent.weight_ih_l0"]
        b_ih, b_hh = sd["a2c_nets.recurrent.bias_ih_l0"], sd["a2c_nets.recurrent.bias_hh_l0"]
        # one step from a zero state: W_hh @ 0 vanishes, only its bias is left
        self.w_ih_t = w_ih.t().contiguous()
        if self.rtype == "GRU":
            # r, z gates: both biases add; n gate: b_in + r * b_hn (torch.nn.GRU gate order r, z, n)
            self.b_rz = (b_ih[: 2 * H] + b_hh[: 2 * H]).contiguous()
            self.b_in, self.b_hn = b_ih[2 * H:].contiguous(), b_hh[2 * H:].contiguous()
        else:
            self.b = (b_ih + b_hh).contiguous()

    @staticmethod
    def _stack(layers, sd, prefix):
        out, idx = [], 0
        for info in layers:
            if info["type"] == "linear":
                w, b = sd[f"{prefix}.{idx}.weight"], sd[f"{prefix}.{idx}.bias"]
                assert w.shape == (info["out_features"], info["in_features"])
                out.append(("linear", w.t().contiguous(), b))
            elif info["type"] == "activation":
                out.append(("act", info.get("activation", "ReLU")))
            else:
                raise ValueError(f"unknown layer type {info['type']}")
            idx += 1
        return out

    @staticmethod
    def _run(stack, x):
        import torch
        for layer in stack:
            if layer[0] == "linear":
                x = torch.addmm(layer[2], x, layer[1])
            else:
                x = _act(layer[1], x)
        return x

    def recurrent_step(self, x):
        """One step of the recurrent layer from a zero state, agents as the batch."""
        import torch
        H = self.hidden_dim
        if self.rtype == "RNN":
            return torch.tanh(torch.addmm(self.b, x, self.w_ih_t))
        if self.rtype == "LSTM":
            g = torch.addmm(self.b, x, self.w_ih_t)                 # gate order i, f, g, o
            c = torch.sigmoid(g[:, :H]) * torch.tanh(g[:, 2 * H:3 * H])
            return torch.sigmoid(g[:, 3 * H:]) * torch.tanh(c)
        g = x @ self.w_ih_t
        rz = torch.sigmoid(g[:, : 2 * H] + self.b_rz)
        n = torch.tanh(g[:, 2 * H:] + self.b_in + rz[:, :H] * self.b_hn)
        return (1.0 - rz[:, H:]) * n

    def forward(self, obs):
        """(n, 69) float32 observations -> (logits (n, 6), value (n, 1))."""
        import torch
        with torch.no_grad():
            shared = self.recurrent_step(self._run(self.feature, obs))
            return self._run(self.actor, shared), self._run(self.critic, shared)

    def logits(self, obs):
        import torch
        with torch.no_grad():
            return self._run(self.actor, self.recurrent_step(self._run(self.feature, obs)))

    @classmethod
    def from_checkpoint(cls, path, device):
        import torch
        ck = torch.load(path, map_location="cpu", weights_only=False)
        return cls(ck["model_config"], ck["model_state_dict"], device)


def load_universe(base_dir, device, metric="latest", num_species=4):
    """checkpoints/universe_<id>/species_<k>/{latest_model,best_<metric>}_epoch_<n>.pt -> [SpeciesPolicy] * 4,
    picking the highest epoch as learn/ckpt.py:36-48 does."""
    out = []
    for k in range(1, num_species + 1):
        pat = "latest_model_epoch_*.pt" if metric == "latest" else f"best_{metric}_epoch_*.pt"
        files = glob.glob(os.path.join(base_dir, f"species_{k}", pat))
        if not files:
            raise FileNotFoundError(f"no {pat} under {base_dir}/species_{k}")
        files.sort(key=lambda p: int(p.split("_")[-1].split(".")[0]), reverse=True)
        out.append(SpeciesPolicy.from_checkpoint(files[0], device))
    return out


class PolicyRunner:
    """Acts for every agent of a simulator, species block by species block:

        runner = PolicyRunner(sim, load_universe("checkpoints/universe_violence", "cuda:0"), seed=0)
        for t in range(T):
            sim.step()
            runner.act()              # pack -> network -> sample -> one-hot rows in sim.action_tensor()
            sim.shift_observations()

    Sampling is deterministic in (seed, act() call number, row): the same rollout replays bit-exactly.
    """

    def __init__(self, sim, policies, seed=0):
        if len(policies) != 4:
            raise ValueError("one policy per species (4)")
        self.sim, self.policies, self.seed = sim, policies, int(seed)
        self.draw = 0

    def species_blocks(self):
        """[(start, end)] * 4 of the current row set: rows are species-major (src/sim/sim.cpp:1148-1149)."""
        A = self.sim.num_agents()
        base = self.sim.export("DEBUG_ROW_BASE").to_torch()[0].tolist()     # first row of (world 0, species k)
        starts = [int(b) for b in base] + [A]
        return [(starts[k], starts[k + 1]) for k in range(4)]

    def act(self, want_index=False):
        import torch
        picked = []
        for (start, end), pol in zip(self.species_blocks(), self.policies):
            if end <= start:
                picked.append(None)
                continue
            obs = self.sim.pack_observations(start, end, prev=False)
            weights = torch.softmax(pol.logits(obs), dim=1).contiguous()
            picked.append(self.sim.sample_actions(start, end, weights, self.seed, self.draw, want_index=want_index))
        self.draw += 1
        return picked
