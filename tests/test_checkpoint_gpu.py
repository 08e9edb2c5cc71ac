"""State checkpoint / restore: a rollout resumed from a blob must continue
bit-identically to the run it was taken from -- and therefore to the CPU oracle."""
import numpy as np
import pytest

from parity_util import ROW_EXPORTS, WORLD_EXPORTS, AGENT_DEBUG, compare, drive, make_pair, gpu_np

pytestmark = pytest.mark.gpu


def _advance(sims, rng, steps, oracle=None):
    """Same actions / hidden state into every simulator of `sims` (and the oracle), `steps` times."""
    import torch
    from oracle_binding import random_actions
    for _ in range(steps):
        A = sims[0].num_agents()
        acts = random_actions(rng, A, "bits")
        hid = rng.standard_normal((A, 16)).astype(np.float32)
        for g in sims:
            assert g.num_agents() == A
            g.action_tensor(False).to_torch().copy_(torch.from_numpy(acts))
            g.hidden_state_tensor(False).to_torch().copy_(torch.from_numpy(hid))
            g.step()
            g.shift_observations()
        if oracle is not None:
            oracle.tensor("ACTION")[:] = acts
            oracle.tensor("HIDDEN_STATE")[:] = hid
            oracle.step()
            oracle.shift_observations()


def _same(a, b, where):
    assert a.num_agents() == b.num_agents(), where
    for name in WORLD_EXPORTS + AGENT_DEBUG + ROW_EXPORTS + ["DEBUG_FOOD"]:
        np.testing.assert_array_equal(gpu_np(a, name), gpu_np(b, name), err_msg=f"{where}: {name}")
    assert a.episode_stats() == b.episode_stats(), where


@pytest.mark.parametrize("steps_before", [1, 7])
def test_restore_resumes_bit_exactly(cuda_device, tmp_path, steps_before):
    from madrona_bots import SimManager
    o, g = make_pair(40, 5, 32)
    rng = np.random.default_rng(3)
    _advance([g], rng, steps_before, oracle=o)
    path = tmp_path / "sim.ckpt"
    blob = g.save_state(path)
    assert blob[:8].tobytes() == b"MBOTSCK1"
    # a fresh handle with ANOTHER seed: everything that matters must come from the blob
    h = SimManager(0, 40, 999, 32)
    h.load_state(path)
    _same(g, h, "right after restore")
    _advance([g, h], rng, 25, oracle=o)
    _same(g, h, "25 steps after restore")
    compare(o, h, "restored run vs oracle")
    # restoring into the handle that kept running rewinds it
    g.load_state(blob)
    h2 = SimManager(0, 40, 1, 32)
    h2.load_state(blob)
    _same(g, h2, "rewound")
    rng2 = np.random.default_rng(77)
    _advance([g, h2], rng2, 5)
    _same(g, h2, "rewound + 5")


def test_restore_rejects_a_mismatched_handle(cuda_device):
    from madrona_bots import SimManager
    g = SimManager(0, 8, 0, 32)
    blob = g.save_state()
    with pytest.raises(RuntimeError, match="different num_worlds"):
        SimManager(0, 9, 0, 32).load_state(blob)
    with pytest.raises(RuntimeError, match="bad magic"):
        g.load_state(np.zeros(blob.size, dtype=np.uint8))
    with pytest.raises(RuntimeError, match="size mismatch|too short"):
        g.load_state(blob[: blob.size // 2])
    g.step_async()
    with pytest.raises(RuntimeError, match="pending"):
        g.save_state()
    g.sync()
    g.save_state()


def test_trajectory_dump_matches_oracle(cuda_device, tmp_path):
    """Headless replacement of the viewer's data path: what the recorder writes for
    the chosen worlds equals the oracle's state of those worlds, step by step."""
    import torch
    from madrona_bots.trajectory import TrajectoryRecorder
    from oracle_binding import random_actions
    o, g = make_pair(6, 2, 32)
    worlds = [1, 4]
    rec = TrajectoryRecorder(g, worlds=worlds)
    rng = np.random.default_rng(5)
    expect = []
    for _ in range(8):
        A = o.num_agents()
        acts = random_actions(rng, A, "onehot")
        o.tensor("ACTION")[:] = acts
        g.action_tensor(False).to_torch().copy_(torch.from_numpy(acts))
        o.step(); g.step()
        n = rec.record()
        cnt, off = o.tensor("DEBUG_WORLD_COUNT")[:, 0], o.tensor("DEBUG_WORLD_OFFSET")[:, 0]
        idx = np.concatenate([np.arange(off[w], off[w] + cnt[w]) for w in worlds])
        assert n == len(idx)
        rows = o.tensor("SENSOR_INDEX")[idx, 0]
        expect.append({"pose": o.tensor("DEBUG_POSE")[idx].copy(), "row": rows.copy(),
                       "depth": o.tensor("SENSOR_DEPTH")[rows].copy(), "semantic": o.tensor("SENSOR_SEMANTIC")[rows].copy(),
                       "reward": o.tensor("REWARD")[rows, 0].copy(), "health": o.tensor("DEBUG_AGENT_HEALTH")[idx, 0].copy(),
                       "target": o.tensor("DEBUG_AGENT_TARGET")[idx, 0].copy(), "action": o.tensor("ACTION")[rows].copy(),
                       "species": o.tensor("DEBUG_AGENT_SPECIES")[idx, 0].astype(np.int8)})
        o.shift_observations(); g.shift_observations()
    path = tmp_path / "traj.npz"
    rec.save(path)
    z = np.load(path)
    so = z["step_offsets"]
    assert len(so) == 9 and list(z["worlds"]) == worlds and z["species_count"].shape == (8, 2, 4)
    for t, e in enumerate(expect):
        for k, v in e.items():
            np.testing.assert_array_equal(z[k][so[t]:so[t + 1]], v, err_msg=f"step {t}: {k}")
    assert set(np.unique(z["world"])) == set(worlds)


def test_corrupt_checkpoint_is_rejected(cuda_device):
    """A blob whose world counts / rows / targets are out of range must not reach the kernels: load_state
    answers MBOTS_EINVAL and leaves a usable, re-initialised handle."""
    from madrona_bots import SimManager
    W, cap = 8, 128
    g = SimManager(0, W, 5, 32)
    for _ in range(3):
        g.step()
    blob = g.save_state()
    slots = W * cap
    off_count = 80 + slots * (16 + 4 + 1 + 2 + 16 + 4)          # header, pose, health, species, target, scratch, row
    assert np.frombuffer(blob[off_count:off_count + 4 * W].tobytes(), dtype=np.int32).tolist() == \
        gpu_np(g, "DEBUG_WORLD_COUNT")[:, 0].tolist()
    bad = blob.copy()
    bad[off_count:off_count + 4] = np.frombuffer(np.int32(100000).tobytes(), dtype=np.uint8)
    h = SimManager(0, W, 9, 32)
    with pytest.raises(RuntimeError, match="corrupt checkpoint"):
        h.load_state(bad)
    h.step()                                                      # still a working simulator
    assert h.num_agents() >= W * 32
    bad2 = blob.copy()
    off_row = 80 + slots * (16 + 4 + 1 + 2 + 16)
    bad2[off_row:off_row + 4] = np.frombuffer(np.int32(-7).tobytes(), dtype=np.uint8)
    with pytest.raises(RuntimeError, match="corrupt checkpoint"):
        h.load_state(bad2)
    h.load_state(blob)                                            # and the intact blob still restores
    np.testing.assert_array_equal(gpu_np(h, "DEBUG_POSE"), gpu_np(g, "DEBUG_POSE"))
