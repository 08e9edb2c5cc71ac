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


def test_rollout_1024_worlds_1000_steps(cuda_device):
    """BASELINE.json configs[1] / SURVEY.md section 7 as written: 1,024 worlds for 1,000 random-action steps,
    every export compared with the oracle every 50 steps (and after the last)."""
    o, g = make_pair(1024, 0, 32)
    drive(o, g, steps=1000, seed=4321, mode="onehot", check_every=50, write_hidden=False)
    st = o.episode_stats()
    assert st[1] == 1000 and st[10] > 0 and st[11] > 0 and st[12] > 0, st


def test_rollout_2500_worlds_partial_scan_tile(cuda_device):
    """Three tiles of the row scan (1,024 worlds each), the last one partial: the cross-tile prefix of the
    species-major row assignment (the reference's global sort, src/sim/sim.cpp:1148-1149) against the oracle."""
    o, g = make_pair(2500, 17, 32)
    drive(o, g, steps=40, seed=99, mode="bits", check_every=5)


def test_rollout_16384_worlds(cuda_device):
    """BASELINE.json configs[2]: 16,384 worlds on one GPU, 20 steps, every export against the oracle."""
    o, g = make_pair(16384, 0, 32)
    drive(o, g, steps=20, seed=77, mode="onehot", check_every=5)


def test_rollout_65536_worlds(cuda_device):
    """The headline size (65,536 worlds on one GPU), 5 steps, every export against the oracle."""
    o, g = make_pair(65536, 0, 32)
    drive(o, g, steps=5, seed=55, mode="onehot", check_every=5, write_hidden=False)


def test_capacity_overflow(cuda_device):
    """Slot table full: births and respawns beyond the capacity are dropped and counted (D1)."""
    o, g = make_pair(32, 3, 32, cap=32)
    drive(o, g, steps=80, seed=5, mode="bits")
    assert o.episode_stats()[13] > 0, "the test must hit the overflow path"


@pytest.mark.parametrize("init_agents", [1, 3, 6, 13])
def test_ragged_init(cuda_device, init_agents):
    o, g = make_pair(33, 9, init_agents)
    drive(o, g, steps=60, seed=2, mode="bits")


@pytest.mark.parametrize("init_agents,cap", [(60, 64), (56, 96), (44, 64)])
def test_worlds_around_the_fast_sensor_limit(cuda_device, init_agents, cap):
    """The sensor's fast variant (32-bit nearest-hit keys, slot 63 = miss sentinel) takes worlds of up to 63
    agents, the full variant the rest: populations that cross 63 / 64 in both directions."""
    o, g = make_pair(24, 21, init_agents, cap=cap)
    counts = drive(o, g, steps=50, seed=13, mode="bits", check_every=5)
    wc = o.tensor("DEBUG_WORLD_COUNT")[:, 0]
    assert max(counts) > 24 * 40 and wc.max() <= cap


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
    # SimManager.species_blocks(): the exact row range of every species, from the totals the step brought to the host
    blocks = g.species_blocks()
    assert [b[0] for b in blocks] == [int(x) for x in rb[0]] and blocks[3][1] == g.num_agents()
    by_row = species[order]
    for k, (a, b) in enumerate(blocks):
        assert (by_row[a:b] == k + 1).all() and b - a == int((species == k + 1).sum())


class _GpuAdapter:
    """The CUDA simulator behind the host-array surface tests/golden/make_golden.run_rollout drives."""

    def __init__(self, worlds, seed, init, cap, off):
        from madrona_bots import SimManager
        self.sim = SimManager(0, worlds, seed, init, max_agents_per_world=cap, world_id_offset=off)

    def num_agents(self):
        return self.sim.num_agents()

    def tensor(self, name):
        return gpu_np(self.sim, name)

    def write(self, name, value):
        import torch
        if len(value):
            self.sim.export(name).to_torch().copy_(torch.from_numpy(value))

    def step(self):
        self.sim.step()

    def shift_observations(self):
        self.sim.shift_observations()

    def episode_stats(self):
        return np.array(self.sim.episode_stats(), dtype=np.int64)


@pytest.mark.parametrize("name", ["w1", "w24_bits", "w16_offset", "w8_cap32", "w5_init13"])
def test_cuda_path_reproduces_golden_rollouts(cuda_device, name):
    """The committed fixtures (generated through the reference's own rand.hpp / math.hpp, tests/golden/
    make_golden.py) against the CUDA path alone -- no oracle in the loop: agent counts and a CRC32 of every
    export after every step."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    from make_golden import DIGEST_EXPORTS, ROLLOUTS, run_rollout
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rollout_ref.npz"))

    sim, counts, digs = run_rollout(_GpuAdapter, ROLLOUTS[name])
    np.testing.assert_array_equal(counts, gold[name + ".counts"])
    bad = np.argwhere(digs != gold[name + ".digests"])
    assert len(bad) == 0, f"first mismatch: step {bad[0][0]}, export {DIGEST_EXPORTS[bad[0][1]]}"
    np.testing.assert_array_equal(sim.episode_stats(), gold[name + ".stats"])


def test_full_size_properties(cuda_device):
    """BASELINE.json's full per-GPU size (65,536 worlds): size-independent properties instead of the oracle.
    Rows are a permutation, species blocks are contiguous and sized by the scan, population bookkeeping closes,
    shift is an idempotent copy, and the depth pixel-0 stamp equals the row index."""
    import torch
    from madrona_bots import SimManager
    W = 65536
    sim = SimManager(0, W, 0, 32)
    g = torch.Generator(device="cuda").manual_seed(1)
    prev = sim.episode_stats()
    n_prev = sim.num_agents()
    for t in range(12):
        A = sim.num_agents()
        a = sim.action_tensor(False).to_torch()
        a.zero_()
        a.scatter_(1, torch.randint(0, 6, (A, 1), generator=g, device="cuda"), 1)
        sim.step()
        st = sim.episode_stats()
        d = [b - c for b, c in zip(st, prev)]
        assert sim.num_agents() == n_prev - d[11] + d[10] + d[12]
        assert d[13] == 0
        prev, n_prev = st, sim.num_agents()
    A = sim.num_agents()
    rows = sim.sensor_index_tensor().to_torch()[:, 0].long()
    assert torch.equal(torch.sort(rows).values, torch.arange(A, device="cuda"))
    species = sim.export("DEBUG_AGENT_SPECIES").to_torch()[:, 0].clone()
    by_row = torch.empty(A, dtype=torch.int32, device="cuda")
    by_row[rows] = species
    assert bool((by_row[1:] >= by_row[:-1]).all()), "rows must be species-major"
    wc = sim.export("DEBUG_WORLD_COUNT").to_torch()[:, 0]
    assert int(wc.sum()) == A and int(wc.min()) >= 32
    depth = sim.depth_tensor(False).to_torch()
    assert torch.equal(depth[:, 0], (torch.arange(A, device="cuda") & 0xFF).to(torch.uint8))
    sem = sim.semantic_tensor(False).to_torch()
    assert bool(((sem == -1) == (depth == 255))[:, 1:].all())
    assert int(sem.max()) <= 4 and int(sem.min()) >= -1
    pos = sim.position_tensor(False).to_torch()
    assert float(pos.min()) >= 0 and float(pos[:, 0].max()) < 128 and float(pos[:, 1].max()) < 96   # respawns are clamped from their first step on
    sim.shift_observations()
    p1 = sim.position_tensor(True).to_torch().clone()
    assert torch.equal(p1, pos)
    assert torch.equal(sim.hidden_state_tensor(True).to_torch(), sim.hidden_state_tensor(False).to_torch())
    st_c, st_p = sim.stats_tensor(False).to_torch(), sim.stats_tensor(True).to_torch()
    assert torch.equal(st_p[:, 1], st_c[:, 0]) and torch.equal(st_p[:, 2:], st_c[:, 2:])
    sim.shift_observations()
    assert torch.equal(sim.position_tensor(True).to_torch(), p1)


def test_reference_plumbing_single_world_1000_steps(cuda_device):
    """BASELINE.json configs[0]: the loop of learn/env.py:46-98 / learn/training_loop.py:36-137 against one world
    for 1,000 steps -- step(), species slices from species_count_tensor().sum(0).cumsum(0), construct_obs of cur
    and prev rows, one-hot actions written per species slice, shift_observations() once per species -- with the
    oracle fed the same action tensor. Every export stays bit-identical."""
    import torch
    o, g = make_pair(1, 69, 32)
    gen = torch.Generator(device="cuda").manual_seed(7)
    for epoch in range(1, 1001):
        o.step(); g.step()
        counts = g.species_count_tensor().to_torch()
        assert counts.shape == (1, 4) and counts.dtype == torch.int32
        ends = counts.sum(dim=0).cumsum(dim=0, dtype=int)
        starts = torch.cat((torch.zeros(1, dtype=torch.int, device="cuda"), ends[:-1]))
        action = g.action_tensor(False).to_torch()
        A = action.shape[0]
        for sp in range(4):
            a, b = int(starts[sp]), min(int(ends[sp]), A)       # counts are taken before the respawn (sim.cpp:804-836)
            obs = torch.cat((g.depth_tensor(False).to_torch()[a:b, :], g.health_tensor(False).to_torch()[a:b, :],
                             g.position_tensor(False).to_torch()[a:b, :], g.semantic_tensor(False).to_torch()[a:b, :],
                             g.surrounding_tensor(False).to_torch()[a:b, :]), dim=1)
            assert obs.shape == (b - a, 69)
            picks = torch.randint(0, 6, (b - a,), generator=gen, device="cuda")
            one_hot = torch.zeros(b - a, 6, device="cuda")
            one_hot.scatter_(1, picks.unsqueeze(1), 1)
            g.shift_observations()                              # inside the species loop, as the reference has it
            action[a:b, :] = one_hot.int()
            o.shift_observations()
            o.tensor("ACTION")[a:b, :] = one_hot.int().cpu().numpy()
        if epoch % 100 == 0 or epoch == 1:
            compare(o, g, f"epoch {epoch}")
    st = o.episode_stats()
    assert st[1] == 1000 and st[11] > 0 and st[12] > 0


def test_async_step_shift_then_sync_matches_oracle(cuda_device):
    """step_async + shift_observations_async + sync (mbots_sync waits for the step only, the shift keeps running
    behind it) gives the same rows as the blocking calls, with and without kernel timing enabled."""
    import torch
    from oracle_binding import random_actions
    o, g = make_pair(96, 21, 32)
    rng = np.random.default_rng(5)
    for t in range(40):
        g.profile(t % 2 == 0)
        A = o.num_agents()
        assert g.num_agents() == A
        acts = random_actions(rng, A, "onehot")
        o.tensor("ACTION")[:] = acts
        g.action_tensor(False).to_torch().copy_(torch.from_numpy(acts))
        o.step(); o.shift_observations()
        g.step_async(); g.shift_observations_async(); g.sync()
        torch.cuda.synchronize()
        compare(o, g, f"async step {t}")
    g.profile(True)
    g.step_async(); g.shift_observations_async(); g.sync()
    ms = g.profile_read()
    assert all(v["calls"] >= 1 for k, v in ms.items() if k in ("step", "export_sense")), ms


@pytest.mark.parametrize("worlds,depth", [(96, 1), (96, 6), (2500, 4)])
def test_rollout_step_graph_and_run_ahead_match_oracle(cuda_device, worlds, depth):
    """mbots_rollout_step_async: action write (row count resolved on the device) + step + shift as one CUDA graph
    launch, with up to `depth` steps enqueued before the host waits (launch geometry then comes from stale totals
    plus slack). The oracle, fed the same action rows, must agree on every export whenever the host catches up."""
    import torch
    from oracle_binding import random_actions
    o, g = make_pair(worlds, 33, 32)
    rng = np.random.default_rng(9)
    rows_cap = worlds * 64
    t = 0
    for rnd in range(8):
        bufs = []
        for k in range(depth):
            A = o.num_agents()
            acts = random_actions(rng, A, "bits" if rnd % 2 else "onehot")
            o.tensor("ACTION")[:] = acts
            o.step(); o.shift_observations()
            dev = torch.full((rows_cap, 6), 7, dtype=torch.int32, device="cuda")     # rows beyond the live count are ignored
            dev[:A] = torch.from_numpy(acts).cuda()
            bufs.append(dev)                                                          # stays alive until the sync below
            g.rollout_step_async(dev, shift=True)
            t += 1
        g.sync()
        torch.cuda.synchronize()
        compare(o, g, f"graph step {t} (depth {depth})")
    import os
    assert g.graph_launch_count() == (0 if os.environ.get("MBOTS_GRAPHS") == "0" else t)
    st = o.episode_stats()
    assert st[10] > 0 and st[11] > 0


def test_graphs_and_plain_launches_agree(cuda_device):
    """The graph path (default), per-kernel timing (plain launches with events) and the blocking step() interleave
    freely on one handle."""
    import torch
    from oracle_binding import random_actions
    o, g = make_pair(64, 2, 32)
    rng = np.random.default_rng(4)
    for t in range(30):
        A = o.num_agents()
        acts = random_actions(rng, A, "onehot")
        o.tensor("ACTION")[:] = acts
        o.step(); o.shift_observations()
        g.profile(t % 3 == 1)
        dev = torch.from_numpy(acts).cuda()
        if t % 3 == 2:
            g.write_actions(dev); g.step(); g.shift_observations()
        else:
            g.rollout_step_async(dev, shift=True); g.sync()
        torch.cuda.synchronize()
        compare(o, g, f"mixed step {t}")


class _GpuRefAdapter:
    """One world on the CUDA path behind the surface tests/golden/make_refsim_golden.run_case drives."""

    def __init__(self, seed, world_id, init):
        from madrona_bots import SimManager
        self.sim = SimManager(0, 1, seed, init, world_id_offset=world_id)

    def num_agents(self):
        return self.sim.num_agents()

    def tensor(self, name):
        return gpu_np(self.sim, name)

    def write(self, name, value):
        import torch
        if len(value):
            self.sim.export(name).to_torch().copy_(torch.from_numpy(value))

    def step(self):
        self.sim.step()

    def shift_observations(self):
        self.sim.shift_observations()


@pytest.mark.parametrize("name", ["s0_w0", "s7_w3_bits", "s21_w4099", "s9_w1_init13", "s5_w2_init48"])
def test_cuda_path_reproduces_reference_text_rollouts(cuda_device, name):
    """tests/golden/refsim_rollouts.npz holds per-step digests of every export of worlds run from the REFERENCE'S
    OWN SOURCE TEXT (src/sim/sim.cpp, M/src/render/ecs_system.cpp, M/src/mw/device/bvh_raycast.cpp under the host
    mock of oracle/refmock/; generator: tests/golden/make_refsim_golden.py). The CUDA path alone -- no oracle, no
    reference on this box -- must reproduce them bit for bit."""
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    from make_refsim_golden import CASES, run_case
    from refsim_binding import REF_EXPORTS
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "refsim_rollouts.npz"))
    sim, counts, digs = run_case(_GpuRefAdapter, CASES[name])
    np.testing.assert_array_equal(counts, gold[name + ".counts"])
    bad = np.argwhere(digs != gold[name + ".digests"])
    assert len(bad) == 0, f"first mismatch: step {bad[0][0]}, export {REF_EXPORTS[bad[0][1]]}"


def test_write_through_a_stale_view_is_detected(cuda_device):
    """include/mbots.h "Pointer lifetime": row exports ping-pong between two row sets, so a view fetched before a step
    points at the dead set afterwards. Debug mode poisons the dead set's writable columns and catches the write."""
    import torch
    from madrona_bots import SimManager
    sim = SimManager(0, 16, 3, 32)
    sim.debug_stale_views(True)
    sim.step()
    for _ in range(3):                                   # the reference's pattern: re-fetch after every step -- never trips
        a = sim.action_tensor(False).to_torch()
        a.zero_(); a[:, 0] = 1
        sim.hidden_state_tensor(False).to_torch().fill_(0.5)
        sim.step()
    stale = sim.action_tensor(False).to_torch()          # fetched now ...
    sim.step()                                           # ... dead after this step
    assert stale.data_ptr() != sim.action_tensor(False).to_torch().data_ptr()
    stale[:4, 1] = 1                                     # lost write
    with pytest.raises(RuntimeError, match="stale|before the previous step"):
        sim.step()
    sim.step()                                           # the handle stays usable
    sim.debug_stale_views(False)
    o, g = make_pair(16, 3, 32)                          # and the mode does not change results
    g.debug_stale_views(True)
    drive(o, g, steps=12, seed=1, mode="bits")


def test_bulk_copy_shift_equals_the_columns_it_copies(cuda_device):
    """shift_observations() runs on the bulk-copy engine (csrc/shift_bulk.cu): 16 KB tiles through a four-stage shared-memory
    ring per CTA. At 12,000 worlds every CTA takes ~8 tiles, so the ring and the mbarrier phases wrap (the small oracle cases
    take one or two tiles per CTA); an odd row count leaves 8- and 4-byte tails behind the 16-byte bulk part. prev := cur on
    every column, with the reference's stats field mix-up (U7)."""
    import torch
    from madrona_bots import SimManager
    sim = SimManager(0, 12000, 3, 32)
    rng = np.random.default_rng(1)
    for t in range(3):
        A = sim.num_agents()
        sim.action_tensor(False).to_torch().copy_(torch.from_numpy(rng.integers(0, 2, (A, 6)).astype(np.int32)))
        sim.hidden_state_tensor(False).to_torch().copy_(torch.from_numpy(rng.standard_normal((A, 16)).astype(np.float32)))
        sim.step()
        cur = {n: gpu_np(sim, n).copy() for n in ("HIDDEN_STATE", "ACTION", "STATS", "POSITION", "SURROUNDING", "HEALTH", "REWARD")}
        untouched = {n: gpu_np(sim, n).copy() for n in ("PREV_SENSOR_DEPTH", "PREV_SENSOR_SEMANTIC", "SENSOR_DEPTH")}
        sim.shift_observations()
        for n, v in cur.items():
            want = v.copy()
            if n == "STATS":
                want[:, 1] = v[:, 0]
            np.testing.assert_array_equal(gpu_np(sim, "PREV_" + n), want, err_msg=f"step {t}: PREV_{n}")
            np.testing.assert_array_equal(gpu_np(sim, n), v, err_msg=f"step {t}: {n} changed")
        for n, v in untouched.items():
            np.testing.assert_array_equal(gpu_np(sim, n), v, err_msg=f"step {t}: {n} changed")
    assert sim.num_agents() * 132 > 444 * 5 * 16384, "too few tiles per CTA to wrap the ring"
