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

__all__ = ["SpeciesPolicy", "FusedPolicy", "PolicyRunner", "load_universe", "load_universe_fixture"]

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

    def __init__(self, config, state_dict, device, dtype="fp32"):
        """dtype: "fp32" (default; matches torch's fp32 modules to 1e-5), "tf32" (fp32 storage, tensor-core
        GEMMs) or "bf16" (weights and activations in bfloat16, fp32 accumulation, fp32 logits)."""
        import torch
        if dtype not in ("fp32", "tf32", "bf16"):
            raise ValueError("dtype must be fp32, tf32 or bf16")
        self.config = config
        self.dtype = dtype
        self.device = torch.device(device)
        wt = torch.bfloat16 if dtype == "bf16" else torch.float32
        self._wt = wt
        sd = {k: v.detach().to(self.device, wt).contiguous() for k, v in state_dict.items()}
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

    def _ctx(self):
        import contextlib
        import torch
        if self.dtype != "tf32":
            return contextlib.nullcontext()

        @contextlib.contextmanager
        def tf32():
            old = torch.backends.cuda.matmul.allow_tf32
            torch.backends.cuda.matmul.allow_tf32 = True
            try:
                yield
            finally:
                torch.backends.cuda.matmul.allow_tf32 = old
        return tf32()

    def forward(self, obs):
        """(n, 69) float32 observations -> (logits (n, 6), value (n, 1)), both float32."""
        import torch
        with torch.no_grad(), self._ctx():
            shared = self.recurrent_step(self._run(self.feature, obs.to(self._wt)))
            return self._run(self.actor, shared).float(), self._run(self.critic, shared).float()

    def logits(self, obs):
        import torch
        with torch.no_grad(), self._ctx():
            return self._run(self.actor, self.recurrent_step(self._run(self.feature, obs.to(self._wt)))).float()

    @classmethod
    def from_checkpoint(cls, path, device, dtype="fp32"):
        import torch
        ck = torch.load(path, map_location="cpu", weights_only=False)
        return cls(ck["model_config"], ck["model_state_dict"], device, dtype)


class FusedPolicy:
    """A SpeciesPolicy compiled for `SimManager.policy_act` (libmbots.so `mbots_policy_act`): the network
    flattened into the program + bf16 weight blob of include/mbots_policy.h, so that observation packing, every
    layer, softmax, sampling and the action write run as ONE tensor-core kernel per species.
    Restrictions of the kernel: hidden width 128, observation width 69, 6 actions (the reference's defaults,
    learn/training_loop.py:228-232) and the actor head Linear-ReLU-Linear of learn/models.py:33-37."""

    def __init__(self, policy, layout="sw128"):
        """layout="sw128": 32 KB pre-swizzled weight blocks for the tcgen05 kernel (policy_umma.cu, the default);
        layout="rowpad": padded rows for the mma.sync kernel (policy_fused.cu)."""
        import numpy as np
        import torch
        from . import _native
        H, PAD = 128, 8
        if layout not in ("sw128", "rowpad"):
            raise ValueError("layout must be 'sw128' or 'rowpad'")
        self.layout = layout
        if policy.hidden_dim != H:
            raise ValueError("the fused kernel is built for hidden_dim 128")
        chunks, size = [], 0

        def put(arr):
            nonlocal size
            ofs = size
            raw = arr.tobytes()
            raw += b"\0" * (-len(raw) % 16)
            chunks.append(raw)
            size += len(raw)
            return ofs

        def bf16_bits(t):      # (rows, k) float tensor -> uint16 bit patterns, round to nearest even
            return t.to(torch.bfloat16).view(torch.int16).cpu().numpy().view(np.uint16)

        def block(w, rows, k_pad):     # w: (out, in) -> one weight block of include/mbots_policy.h, zero padded
            if layout == "rowpad":     # [rows][k_pad + 8] bf16
                out = np.zeros((rows, k_pad + PAD), dtype=np.uint16)
                out[: w.shape[0], : w.shape[1]] = bf16_bits(w)
                return out
            dense = np.zeros((H, H), dtype=np.uint16)
            dense[: w.shape[0], : w.shape[1]] = bf16_bits(w)
            row, k = np.meshgrid(np.arange(H), np.arange(H), indexing="ij")
            ofs = (k // 64) * 16384 + (row // 8) * 1024 + (row % 8) * 128 + ((((k % 64) // 8) ^ (row % 8)) * 16) + (k % 8) * 2
            out = np.zeros(H * H, dtype=np.uint16)      # K-major SWIZZLE_128B image, 32 KB
            out[ofs.ravel() // 2] = dense.ravel()
            return out

        def f32(t):
            return t.detach().float().cpu().numpy().astype(np.float32)

        ops = []

        def linear(w_t, b, act, first):
            w = w_t.t().float()                         # stored transposed in SpeciesPolicy
            k_pad = 80 if first else H
            if w.shape[0] != H or w.shape[1] > k_pad or (not first and w.shape[1] != H):
                raise ValueError("the fused kernel needs 69 -> 128 -> ... -> 128 linear layers")
            ops.append((_native.OP_LINEAR, act, k_pad, put(block(w, H, k_pad)), put(f32(b))))

        def stack(layers, first):
            i = 0
            while i < len(layers):
                kind = layers[i]
                if kind[0] != "linear":
                    raise ValueError("an activation without a preceding linear layer")
                act = "NONE"
                if i + 1 < len(layers) and layers[i + 1][0] == "act":
                    act = layers[i + 1][1]
                    i += 1
                if i + 1 < len(layers) and layers[i + 1][0] == "act":
                    raise ValueError("two activations in a row")
                linear(kind[1], kind[2], _native.ACT[act], first)
                first = False
                i += 1

        stack(policy.feature, True)
        w_ih = policy.w_ih_t.t().float()                # (gates * H, H)
        if w_ih.shape[1] != H:
            raise ValueError("recurrent input width must be 128")
        if policy.rtype == "RNN":
            ops.append((_native.OP_LINEAR, _native.ACT["Tanh"], H, put(block(w_ih, H, H)), put(f32(policy.b))))
        elif policy.rtype == "LSTM":                    # torch gate order i, f, g, o; f is unused from c0 = 0
            b = policy.b.float()
            w_ofs = put(np.concatenate([block(w_ih[j * H:(j + 1) * H], H, H) for j in (0, 2, 3)]))
            ops.append((_native.OP_LSTM0, 0, H, w_ofs, put(np.concatenate([f32(b[j * H:(j + 1) * H]) for j in (0, 2, 3)]))))
        else:                                           # GRU: r, z, n
            w_ofs = put(np.concatenate([block(w_ih[j * H:(j + 1) * H], H, H) for j in range(3)]))
            b_rz = policy.b_rz.float()
            ops.append((_native.OP_GRU0, 0, H, w_ofs,
                        put(np.concatenate([f32(b_rz[:H]), f32(b_rz[H:]), f32(policy.b_in), f32(policy.b_hn)]))))
        actor = policy.actor
        if [l[0] for l in actor] != ["linear", "act", "linear"] or actor[1][1] != "ReLU":
            raise ValueError("actor head must be Linear-ReLU-Linear")
        linear(actor[0][1], actor[0][2], _native.ACT["ReLU"], False)
        w_out, b_out = actor[2][1].t().float(), actor[2][2]
        if w_out.shape != (6, H):
            raise ValueError("the head must map 128 -> 6")
        bias8 = np.zeros(8, dtype=np.float32)
        bias8[:6] = f32(b_out)
        ops.append((_native.OP_HEAD, 0, H, put(block(w_out, 8, H)), put(bias8)))
        if len(ops) > 12:
            raise ValueError("network too deep for the fused kernel (12 ops)")
        self.program = _native.PolicyProgram()
        self.program.n_ops = len(ops)
        self.program.layout = _native.LAYOUT_SW128 if layout == "sw128" else _native.LAYOUT_ROWPAD
        for i, (op, act, k_pad, w_ofs, b_ofs) in enumerate(ops):
            self.program.ops[i] = _native.PolicyOp(op, act, k_pad, w_ofs, b_ofs)
        host = np.frombuffer(b"".join(chunks), dtype=np.uint8).copy()
        self.blob = torch.from_numpy(host).to(policy.device)
        self.policy = policy


def load_universe(base_dir, device, metric="latest", num_species=4, dtype="fp32"):
    """checkpoints/universe_<id>/species_<k>/{latest_model,best_<metric>}_epoch_<n>.pt -> [SpeciesPolicy] * 4,
    picking the highest epoch as learn/ckpt.py:36-48 does."""
    out = []
    for k in range(1, num_species + 1):
        pat = "latest_model_epoch_*.pt" if metric == "latest" else f"best_{metric}_epoch_*.pt"
        files = glob.glob(os.path.join(base_dir, f"species_{k}", pat))
        if not files:
            raise FileNotFoundError(f"no {pat} under {base_dir}/species_{k}")
        files.sort(key=lambda p: int(p.split("_")[-1].split(".")[0]), reverse=True)
        out.append(SpeciesPolicy.from_checkpoint(files[0], device, dtype))
    return out


def load_universe_fixture(npz_path, device, dtype="fp32", num_species=4):
    """The four species networks of checkpoints/universe_violence from the fixture tests/golden/make_policy_golden.py
    writes (weights + model_config per species, fp32): what a box without /root/reference loads instead."""
    import json
    import numpy as np
    import torch
    z = np.load(npz_path)
    out = []
    for k in range(1, num_species + 1):
        cfg = json.loads(bytes(z[f"sp{k}.config_json"]).decode())
        pre = f"sp{k}.w."
        sd = {n[len(pre):]: torch.from_numpy(z[n]) for n in z.files if n.startswith(pre)}
        out.append(SpeciesPolicy(cfg, sd, device, dtype))
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

    def __init__(self, sim, policies, seed=0, fused=False):
        """fused=True (or "sw128" / "rowpad", the FusedPolicy layout): each species acts through one
        `mbots_policy_act` kernel (bf16 tensor cores) instead of pack kernel + library GEMMs + sample kernel."""
        if len(policies) != 4:
            raise ValueError("one policy per species (4)")
        self.sim, self.policies, self.seed = sim, policies, int(seed)
        layout = fused if isinstance(fused, str) else "sw128"
        self.fused = [FusedPolicy(p, layout) if not isinstance(p, FusedPolicy) else p for p in policies] if fused else None
        self.draw = 0

    def species_blocks(self):
        """[(start, end)] * 4 of the current row set: rows are species-major (src/sim/sim.cpp:1148-1149)."""
        return self.sim.species_blocks()

    def act(self, want_index=False):
        import torch
        picked = []
        for k, ((start, end), pol) in enumerate(zip(self.species_blocks(), self.policies)):
            if end <= start:
                picked.append(None)
                continue
            if self.fused is not None:
                picked.append(self.sim.policy_act(start, end, self.fused[k], self.seed, self.draw, want_index=want_index))
                continue
            obs = self.sim.pack_observations(start, end, prev=False)
            weights = torch.softmax(pol.logits(obs), dim=1).contiguous()
            picked.append(self.sim.sample_actions(start, end, weights, self.seed, self.draw, want_index=want_index))
        self.draw += 1
        return picked

    def act_async(self, out_index=None, out_logits=None, one_launch=True):
        """act() without any host wait: ONE kernel for the four species (or four with one_launch=False) whose row ranges
        are resolved on the device (fused SW128 policies only). With `sim.rollout_step_async(None, shift=True)` after it, a
        whole rollout step -- act, step, sensor, shift -- is two launches and the host may run several steps ahead; same
        draws as act(), so the same rollout."""
        if self.fused is None:
            raise ValueError("act_async needs fused policies (PolicyRunner(..., fused='sw128'))")
        if one_launch and not getattr(self, "_too_big_for_one_launch", False):
            try:
                self.sim.policy_act_all(self.fused, self.seed, self.draw, out_index=out_index, out_logits=out_logits)
            except RuntimeError as e:
                if "biases exceed" not in str(e):
                    raise
                self._too_big_for_one_launch = True       # four deep networks: one launch per species from now on
                return self.act_async(out_index, out_logits, one_launch=False)
        else:
            for k in range(4):
                self.sim.policy_act_species(k, self.fused[k], self.seed, self.draw, out_index=out_index, out_logits=out_logits)
        self.draw += 1
