"""madrona_bots.policy, host side: the network evaluation built from a reference-format
checkpoint equals torch's own modules run as "one recurrent step per agent from a zero state"."""
import os
import sys

import pytest
import torch
import torch.nn as nn

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "madrona-bots_b200"))
from madrona_bots.policy import SpeciesPolicy  # noqa: E402


class _A2C(nn.Module):            # same attribute names as learn/models.py:6-12, so state_dict keys match
    def __init__(self, feature, recurrent, actor, critic):
        super().__init__()
        self.feature, self.recurrent, self.actor, self.critic = feature, recurrent, actor, critic


class _Net(nn.Module):
    def __init__(self, a2c):
        super().__init__()
        self.a2c_nets = a2c


def _make(rtype, acts, H=128, obs=69):
    layers, cfg = [nn.Linear(obs, H)], [{"type": "linear", "in_features": obs, "out_features": H}]
    for a in acts:
        layers += [nn.Linear(H, H), getattr(nn, a)()]
        cfg += [{"type": "linear", "in_features": H, "out_features": H}, {"type": "activation", "activation": a}]
    head = lambda out: nn.Sequential(nn.Linear(H, H), nn.ReLU(True), nn.Linear(H, out))  # noqa: E731
    head_cfg = lambda out: [{"type": "linear", "in_features": H, "out_features": H},  # noqa: E731
                            {"type": "activation", "activation": "ReLU"},
                            {"type": "linear", "in_features": H, "out_features": out}]
    net = _Net(_A2C(nn.Sequential(*layers), getattr(nn, rtype)(H, H), head(6), head(1)))
    config = {"layers": cfg, "actor": head_cfg(6), "critic": head_cfg(1),
              "recurrent": {"type": rtype, "input_dim": H, "hidden_dim": H}}
    return net, config


@pytest.mark.parametrize("rtype,acts", [("GRU", ["Tanh", "ReLU"]), ("LSTM", ["LeakyReLU", "ReLU"]),
                                        ("RNN", ["LogSigmoid", "Tanh", "ELU"])])
def test_forward_equals_torch_modules(rtype, acts):
    torch.manual_seed(3)
    net, config = _make(rtype, acts)
    pol = SpeciesPolicy(config, net.state_dict(), "cpu")
    x = torch.randn(257, 69) * 3
    with torch.no_grad():
        f = net.a2c_nets.feature(x)
        shared, _ = net.a2c_nets.recurrent(f.unsqueeze(0))          # (seq 1, batch n, H), zero initial state
        shared = shared[0]
        ref_logits, ref_value = net.a2c_nets.actor(shared), net.a2c_nets.critic(shared)
    logits, value = pol.forward(x)
    torch.testing.assert_close(logits, ref_logits, rtol=1e-5, atol=1e-5)     # fp32, different summation order
    torch.testing.assert_close(value, ref_value, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(pol.logits(x), logits, rtol=0, atol=0)


def test_checkpoint_file_round_trip(tmp_path):
    """The file format of learn/ckpt.py:28-32."""
    net, config = _make("GRU", ["Tanh"])
    path = tmp_path / "latest_model_epoch_7.pt"
    torch.save({"model_state_dict": net.state_dict(), "optimizer_state_dict": {}, "model_config": config}, path)
    pol = SpeciesPolicy.from_checkpoint(str(path), "cpu")
    assert pol.rtype == "GRU" and pol.logits(torch.zeros(2, 69)).shape == (2, 6)


def test_unknown_layers_are_rejected():
    net, config = _make("RNN", ["Tanh"])
    bad = dict(config, recurrent={"type": "Transformer", "input_dim": 128, "hidden_dim": 128})
    with pytest.raises(ValueError):
        SpeciesPolicy(bad, net.state_dict(), "cpu")


def test_species_policy_equals_reference_models_py_on_shipped_weights():
    """tests/golden/policy_universe_violence.npz: logits and values of the reference's OWN learn/models.py
    (ActorCritic.forward, one agent at a time = one recurrent step from a zero state, DEVIATIONS.md D4) on the shipped
    checkpoints/universe_violence weights, 512 fixed observation rows. The host-side evaluation must agree in fp32."""
    import numpy as np
    from madrona_bots.policy import load_universe_fixture
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "policy_universe_violence.npz")
    z = np.load(path)
    obs = torch.from_numpy(np.concatenate([z["depth"].astype(np.float32), z["health"], z["position"],
                                           z["semantic"].astype(np.float32), z["surrounding"]], axis=1))
    pols = load_universe_fixture(path, "cpu")
    kinds = [p.rtype for p in pols]
    assert kinds == ["GRU", "LSTM", "RNN", "RNN"], kinds
    for k, pol in enumerate(pols, 1):
        logits, value = pol.forward(obs)
        ref_l, ref_v = torch.from_numpy(z[f"sp{k}.logits"]), torch.from_numpy(z[f"sp{k}.values"])
        scale = float(ref_l.abs().max())
        assert float((logits - ref_l).abs().max()) <= 2e-5 * scale, (k, float((logits - ref_l).abs().max()), scale)
        assert float((value - ref_v).abs().max()) <= 2e-5 * float(ref_v.abs().max()) + 1e-4
        # the reference's loop feeds all agents as ONE sequence (learn/models.py:163-167): its row 0 is the zero-state step
        np.testing.assert_allclose(z[f"sp{k}.logits_as_one_sequence_row0"], z[f"sp{k}.logits"][0], rtol=1e-5, atol=1e-3)
