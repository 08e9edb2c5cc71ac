"""learn-compatible training step: the MemoryA2C interface the reference's loop calls
(learn/training_loop.py:57-60,89-95) and `train_epoch` driven over the CPU oracle through the SimManager surface."""
import numpy as np
import pytest
import torch

from oracle_binding import OracleSim, oracle_lib

pytestmark = pytest.mark.skipif(oracle_lib() is None, reason="oracle library not built")


class _T:
    def __init__(self, a):
        self.a = a

    def to_torch(self):
        return torch.from_numpy(self.a)         # shares the oracle's memory: writes land in the simulator


class OracleManager:
    """The reference's SimManager surface (src/entry/bindings.cpp) over the CPU oracle."""

    def __init__(self, *a):
        self.o = OracleSim(*a)

    def step(self):
        self.o.step()

    def shift_observations(self):
        self.o.shift_observations()

    def species_count_tensor(self):
        return _T(self.o.tensor("SPECIES_COUNT"))

    def species_blocks(self):
        starts = [int(b) for b in self.o.tensor("DEBUG_ROW_BASE")[0]] + [self.o.num_agents()]
        return [(starts[k], starts[k + 1]) for k in range(4)]


for _name, _cur, _prev in (("depth", "SENSOR_DEPTH", "PREV_SENSOR_DEPTH"), ("semantic", "SENSOR_SEMANTIC", "PREV_SENSOR_SEMANTIC"),
                           ("reward", "REWARD", "PREV_REWARD"), ("position", "POSITION", "PREV_POSITION"),
                           ("health", "HEALTH", "PREV_HEALTH"), ("surrounding", "SURROUNDING", "PREV_SURROUNDING"),
                           ("action", "ACTION", "PREV_ACTION"), ("hidden_state", "HIDDEN_STATE", "PREV_HIDDEN_STATE")):
    setattr(OracleManager, _name + "_tensor",
            (lambda c, p: lambda self, is_prev=False: _T(self.o.tensor(p if is_prev else c)))(_cur, _prev))


@pytest.mark.parametrize("rtype", ["GRU", "LSTM", "RNN"])
def test_memory_a2c_interface(rtype):
    from madrona_bots.learn import MemoryA2C
    torch.manual_seed(0)
    m = MemoryA2C(rtype=rtype, activations=("LogSigmoid", "Tanh", "ELU"))
    obs, mem = torch.randn(7, 69), torch.randn(7, 16)
    logits, values, sm = m.forward(obs, mem)
    assert logits.shape == (7, 6) and values.shape == (7, 1) and sm.shape == (7, 128)
    nm = m.generate_memory(obs, sm)
    assert nm.shape == (7, 16) and not nm.requires_grad and float(nm.abs().max()) <= 1.0
    # the memory matters, and agents are a batch (row i depends on row i only)
    l2, _, _ = m.forward(obs, mem + 1.0)
    assert not torch.allclose(l2, logits)
    l3, _, _ = m.forward(obs[2:5], mem[2:5])
    assert torch.allclose(l3, logits[2:5], atol=1e-6)
    # compute_loss is the reference's formula (learn/models.py:181-184)
    lp, r, pv, nv = torch.randn(7), torch.randn(7), torch.randn(7), torch.randn(7)
    a, c = MemoryA2C.compute_loss(lp, r, pv, nv)
    assert torch.allclose(a, -(lp * (r + nv - pv)).sum()) and torch.allclose(c, torch.nn.functional.smooth_l1_loss(r, pv))


def test_train_epoch_over_the_oracle():
    from madrona_bots.learn import MemoryA2C, construct_obs, train_epoch
    torch.manual_seed(1)
    sim = OracleManager(3, 7, 32)
    nets = [MemoryA2C(rtype=t) for t in ("GRU", "LSTM", "RNN", "GRU")]
    opts = [torch.optim.Adam(n.parameters(), lr=1e-3) for n in nets]
    before = [p.detach().clone() for p in nets[0].parameters()]
    g = torch.Generator().manual_seed(5)
    for epoch in range(4):
        res = train_epoch(sim, nets, opts, generator=g)
        A = sim.o.num_agents()
        assert len(res) == 4 and sum(r[2] for r in res) == A
        assert all(np.isfinite(r[0]) and np.isfinite(r[1]) for r in res)
        act = sim.o.tensor("ACTION")
        assert act.shape == (A, 6) and (act.sum(axis=1) == 1).all()          # every agent got a one-hot action
        mem = sim.o.tensor("HIDDEN_STATE")
        assert np.abs(mem).max() <= 1.0 and np.abs(mem).max() > 0.0          # memories were written back
        # the loop shifts once per species; after the last shift prev == cur for the shifted columns
        assert np.array_equal(sim.o.tensor("PREV_POSITION"), sim.o.tensor("POSITION"))
    assert any(not torch.equal(a, b) for a, b in zip(before, nets[0].parameters()))   # the optimiser stepped
    obs = construct_obs(sim, 0, 5)
    assert obs.shape == (5, 69) and obs.dtype == torch.float32
    assert torch.equal(obs[:, 33:35], torch.from_numpy(sim.o.tensor("POSITION")[:5]))
