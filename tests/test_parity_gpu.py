"""GPU parity tests proper: the CUDA path (through the C-ABI / madrona_bots.SimManager) against
the CPU oracle on the same seeded inputs. Bit-exact on every export (see parity_util.FLOAT_TOL)."""
import numpy as np
import pytest

from parity_util import compare, drive, make_pair, gpu_np

pytestmark = pytest.mark.gpu


def test_init_matches_oracle(cuda_device):
    o, g = make_pair(64, 0, 32)
    compare(o, g, "init")
    assert g.num_agents() == 64 * 32


@pytest.mark.parametrize("mode", ["onehot", "bits"])
def test_rollout_small(cuda_device, mode):
    o, g = make_pair(48, 7, 32)
    counts = drive(o, g, steps=120, seed=11, mode=mode)
    st = o.episode_stats()
    # the rollout must actually exercise the ragged paths it claims to cover
    assert st[10] > 0 and st[11] > 0 and st[12] > 0 and st[14] >= 0, st
    assert len(set(counts)) > 1


def test_rollout_1024_worlds(cuda_device):
    """BASELINE.json configs[1]: numWorlds=1,024 random-action rollout, bit-exact vs the oracle."""
    o, g = make_pair(1024, 0, 32)
    drive(o, g, steps=60, seed=1234, mode="onehot", check_every=10)


def test_capacity_overflow(cuda_device):
    """Slot table full: births and respawns beyond the capacity are dropped and counted (D1)."""
    o, g = make_pair(32, 3, 32, cap=32)
    drive(o, g, steps=80, seed=5, mode="bits")
    assert o.episode_stats()[13] > 0, "the test must hit the overflow path"


@pytest.mark.parametrize("init_agents", [1, 3, 6, 13])
def test_ragged_init(cuda_device, init_agents):
    o, g = make_pair(33, 9, init_agents)
    drive(o, g, steps=60, seed=2, mode="bits")


def test_zero_agents(cuda_device):
    o, g = make_pair(8, 1, 0)
    drive(o, g, steps=5, seed=2)
    assert g.num_agents() == 0
    assert g.action_tensor(False).to_torch().shape == (0, 6)


def test_large_capacity(cuda_device):
    o, g = make_pair(16, 4, 200, cap=512)
    drive(o, g, steps=40, seed=8, mode="bits", check_every=5)


def test_world_id_offset_shards_match(cuda_device):
    """Worlds are independent and keyed on their GLOBAL id: simulating worlds [8, 16) alone gives the
    same per-world state as simulating [0, 16) (multi-GPU sharding is result-invariant)."""
    import torch
    from madrona_bots import SimManager
    full = SimManager(0, 16, 21, 32)
    part = SimManager(0, 8, 21, 32, world_id_offset=8)
    for t in range(40):
        for sim in (full, part):
            a = sim.action_tensor(False).to_torch()
            # an action that depends only on the agent's own position, so both runs agree on it
            pos = sim.position_tensor(False).to_torch()
            idx = ((pos[:, 0] * 7.0 + pos[:, 1] * 3.0).long() + t) % 6
            a.zero_()
            a.scatter_(1, idx.unsqueeze(1), 1)
            sim.step()
    wc_f, wc_p = gpu_np(full, "DEBUG_WORLD_COUNT")[8:], gpu_np(part, "DEBUG_WORLD_COUNT")
    np.testing.assert_array_equal(wc_f, wc_p)
    off = gpu_np(full, "DEBUG_WORLD_OFFSET")[8, 0]
    np.testing.assert_array_equal(gpu_np(full, "DEBUG_POSE")[off:], gpu_np(part, "DEBUG_POSE"))
    np.testing.assert_array_equal(gpu_np(full, "DEBUG_AGENT_HEALTH")[off:], gpu_np(part, "DEBUG_AGENT_HEALTH"))
    np.testing.assert_array_equal(gpu_np(full, "SPECIES_COUNT")[8:], gpu_np(part, "SPECIES_COUNT"))


def test_api_surface(cuda_device):
    """Shapes and dtypes of the reference's Python surface (src/entry/mgr.cpp:199-422)."""
    import torch
    from madrona_bots import SimManager
    sim = SimManager(0, 4, 0, 32)
    sim.step()
    A = sim.num_agents()
    spec = {
        "action_tensor": (torch.int32, 6), "hidden_state_tensor": (torch.float32, 16),
        "reward_tensor": (torch.float32, 1), "position_tensor": (torch.float32, 2),
        "health_tensor": (torch.float32, 1), "surrounding_tensor": (torch.float32, 2),
        "stats_tensor": (torch.int32, 4), "semantic_tensor": (torch.int8, 32), "depth_tensor": (torch.uint8, 32),
    }
    for name, (dt, cols) in spec.items():
        for prev in (False, True):
            t = getattr(sim, name)(prev).to_torch()
            assert t.dtype == dt and tuple(t.shape) == (A, cols) and t.is_cuda, (name, prev, t.dtype, t.shape)
    sc = sim.species_count_tensor().to_torch()
    assert sc.dtype == torch.int32 and tuple(sc.shape) == (4, 4)
    # views are writable and zero-copy: a write through one view is seen through another
    a = sim.action_tensor(False).to_torch()
    a[:, :] = 0
    a[0, 4] = 1
    assert int(sim.action_tensor(False).to_torch()[0, 4]) == 1
    sim.set_action(1, 1, 0, 0, 0, 0, 0)
    assert sim.action_tensor(False).to_torch()[1].tolist() == [1, 0, 0, 0, 0, 0]
    assert sim.agent_offset_for_world(0) == 0
    assert sim.agent_offset_for_world(1) == int(gpu_np(sim, "DEBUG_WORLD_COUNT")[0, 0])
    with pytest.raises(RuntimeError):
        sim.agent_offset_for_world(99)


def test_species_blocks_are_contiguous(cuda_device):
    """training_loop.py:43-45 slices species blocks out of the row exports."""
    o, g = make_pair(64, 5, 32)
    drive(o, g, steps=30, seed=3, check_every=30)
    rb = gpu_np(g, "DEBUG_ROW_BASE")
    species = gpu_np(g, "DEBUG_AGENT_SPECIES")[:, 0]
    rows = gpu_np(g, "SENSOR_INDEX")[:, 0]
    order = np.argsort(rows)
    assert np.array_equal(np.sort(rows), np.arange(len(rows)))
    assert np.all(np.diff(species[order]) >= 0), "rows must be species-major"
    assert rb[0, 0] == 0
