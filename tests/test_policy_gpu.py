"""Policy-inference ends on the GPU: the fused construct_obs kernel against torch.cat of the
exports, the sampling kernel against a host restatement built on the same Threefry primitives, and a
policy-in-the-loop rollout that stays bit-exact with the CPU oracle fed the same actions."""
import ctypes

import numpy as np
import pytest

from parity_util import compare, gpu_np, make_pair
from test_policy import _make

pytestmark = pytest.mark.gpu


def _construct_obs(sim, start, end, prev):
    """learn/util.py:23-29, verbatim in structure."""
    import torch
    return torch.cat((sim.depth_tensor(prev).to_torch()[start:end, :],
                      sim.health_tensor(prev).to_torch()[start:end, :],
                      sim.position_tensor(prev).to_torch()[start:end, :],
                      sim.semantic_tensor(prev).to_torch()[start:end, :],
                      sim.surrounding_tensor(prev).to_torch()[start:end, :]), dim=1)


def _random_steps(sim, steps, seed):
    import torch
    g = torch.Generator(device="cuda").manual_seed(seed)
    for _ in range(steps):
        a = sim.action_tensor(False).to_torch()
        a.zero_()
        a.scatter_(1, torch.randint(0, 6, (a.shape[0], 1), generator=g, device="cuda"), 1)
        sim.step()
        sim.shift_observations()
    sim.step()


def test_pack_observations_equals_construct_obs(cuda_device):
    import torch
    from madrona_bots import SimManager
    sim = SimManager(0, 40, 3, 32)
    _random_steps(sim, 6, 1)
    A = sim.num_agents()
    for prev in (False, True):
        for start, end in ((0, A), (17, A - 5), (A // 2, A // 2), (A - 1, A)):
            got = sim.pack_observations(start, end, prev=prev)
            want = _construct_obs(sim, start, end, prev).to(torch.float32)
            assert got.shape == (end - start, 69) and got.dtype == torch.float32
            assert torch.equal(got, want)
    assert float(sim.pack_observations(0, A)[:, 35:67].min()) == -1.0       # a missed ray's semantic
    with pytest.raises(RuntimeError, match="row range"):
        sim.pack_observations(0, A + 1)


def _agent_ids(sim, start, end, world_id_offset=0):
    """(GLOBAL world id << 10) | slot of the agents that own rows [start, end) -- what the sampling is keyed on."""
    o = gpu_np(sim, "DEBUG_ROW_OWNER")[start:end, 0].astype(np.int64) & 0xFFFFFFFF
    return ((world_id_offset + (o >> 10)) << 10) | (o & 1023)


def _host_sample(shim, weights, seed, draw, ids):
    """The kernel's definition, restated on the host with the product's own Threefry (prims shim)."""
    k0, k1, k2 = (ctypes.c_uint32 * 2)(), (ctypes.c_uint32 * 2)(), (ctypes.c_uint32 * 2)()
    shim.shim_init_key(seed, k0)
    shim.shim_split(k0[0], k0[1], draw, 0, k1)
    out = np.empty(len(weights), dtype=np.int32)
    f32 = np.float32
    for r, w in enumerate(weights.astype(np.float32)):
        shim.shim_split(k1[0], k1[1], int(ids[r]) & 0xFFFFFFFF, int(ids[r]) >> 32, k2)
        u = f32(shim.shim_sample_uniform(k2[0], k2[1]))
        total = w[0]
        for k in range(1, 6):
            total = f32(total + w[k])
        thr, c, pick = f32(u * total), f32(0), 5
        for k in range(6):
            c = f32(c + w[k])
            if thr < c:
                pick = k
                break
        out[r] = pick
    return out


def test_sample_actions_matches_host_restatement(cuda_device):
    import torch
    from madrona_bots import SimManager
    from prims_binding import shim
    sim = SimManager(0, 24, 5, 32)
    sim.step()
    A = sim.num_agents()
    start, end = 11, A - 7
    g = torch.Generator(device="cuda").manual_seed(4)
    w = torch.rand((end - start, 6), generator=g, device="cuda")
    w[::7, 2] = 0.0                      # zero-weight actions must never be picked
    w[5] = torch.tensor([0, 0, 0, 0, 0, 1.0])
    before = sim.action_tensor(False).to_torch().clone()
    idx = sim.sample_actions(start, end, w, seed=9, draw=3, want_index=True)
    want = _host_sample(shim(), w.cpu().numpy(), 9, 3, _agent_ids(sim, start, end))
    np.testing.assert_array_equal(idx.cpu().numpy(), want)
    act = sim.action_tensor(False).to_torch()
    assert torch.equal(act[start:end], torch.nn.functional.one_hot(idx.long(), 6).to(torch.int32))
    assert torch.equal(act[:start], before[:start]) and torch.equal(act[end:], before[end:])
    assert int(idx[5]) == 5 and not bool((idx[::7] == 2).any())
    # another draw counter gives another sample; the same one repeats
    idx2 = sim.sample_actions(start, end, w, seed=9, draw=4, want_index=True)
    idx3 = sim.sample_actions(start, end, w, seed=9, draw=3, want_index=True)
    assert not torch.equal(idx, idx2) and torch.equal(idx, idx3)
    # the empirical distribution follows the weights
    big = torch.tensor([[1.0, 2.0, 3.0, 4.0, 0.0, 10.0]], device="cuda").repeat(end - start, 1).contiguous()
    hist = torch.zeros(6)
    for d in range(40):
        hist += torch.bincount(sim.sample_actions(start, end, big, 1, 100 + d, want_index=True).long().cpu(), minlength=6)
    p = hist / hist.sum()
    np.testing.assert_allclose(p.numpy(), np.array([1, 2, 3, 4, 0, 10]) / 20.0, atol=0.01)


def test_policy_in_the_loop_rollout_matches_oracle(cuda_device):
    """Species networks drive the simulation; the oracle is fed the sampled actions and must stay bit-identical."""
    import torch
    from madrona_bots.policy import PolicyRunner, SpeciesPolicy
    torch.manual_seed(0)
    pols = []
    for rtype, acts in (("GRU", ["Tanh", "ReLU"]), ("LSTM", ["LeakyReLU", "ReLU"]), ("RNN", ["LogSigmoid", "Tanh", "ReLU"]),
                        ("RNN", ["LogSigmoid", "Tanh", "LeakyReLU"])):          # the shapes of universe_violence
        net, config = _make(rtype, acts)
        pols.append(SpeciesPolicy(config, net.state_dict(), "cuda:0"))
    o, g = make_pair(32, 6, 32)
    o.step(); g.step()
    runner = PolicyRunner(g, pols, seed=2)
    picked_hist = torch.zeros(6)
    for t in range(12):
        blocks = runner.species_blocks()
        assert blocks[0][0] == 0 and blocks[3][1] == g.num_agents() and all(a <= b for a, b in blocks)
        sp = g.export("DEBUG_AGENT_SPECIES").to_torch()[:, 0].clone()
        rows = g.sensor_index_tensor().to_torch()[:, 0].long()
        by_row = torch.empty_like(sp); by_row[rows] = sp
        for k, (a, b) in enumerate(blocks):
            assert bool((by_row[a:b] == k + 1).all())
        picked = runner.act(want_index=True)
        act = g.action_tensor(False).to_torch()
        assert bool((act.sum(dim=1) == 1).all())
        picked_hist += torch.bincount(torch.cat([p for p in picked if p is not None]).long().cpu(), minlength=6)
        o.tensor("ACTION")[:] = act.cpu().numpy()
        o.step(); g.step()
        compare(o, g, f"policy step {t}")
        o.shift_observations(); g.shift_observations()
    assert int((picked_hist > 0).sum()) >= 2


def _bf16_reference_logits(pol, obs):
    """The fused kernel's arithmetic restated with torch: bf16 weights and layer inputs, fp32 accumulation,
    fp32 bias / activation, GRU helper terms rounded to bf16 as the kernel keeps them."""
    import torch
    import torch.nn.functional as F
    r = lambda t: t.to(torch.bfloat16).float()  # noqa: E731
    acts = {"Tanh": torch.tanh, "ELU": F.elu, "LogSigmoid": F.logsigmoid, "LeakyReLU": F.leaky_relu, "ReLU": F.relu}

    def run(stack, x):
        for layer in stack:
            x = r(x) @ r(layer[1].float()) + layer[2].float() if layer[0] == "linear" else acts[layer[1]](x)
        return x
    H = pol.hidden_dim
    x = run(pol.feature, obs)
    g = r(x) @ r(pol.w_ih_t.float())
    if pol.rtype == "RNN":
        h = torch.tanh(g + pol.b)
    elif pol.rtype == "LSTM":
        g = g + pol.b
        h = torch.sigmoid(g[:, 3 * H:]) * torch.tanh(torch.sigmoid(g[:, :H]) * torch.tanh(g[:, 2 * H:3 * H]))
    else:
        rz = torch.sigmoid(g[:, :2 * H] + pol.b_rz)
        h = r(1.0 - rz[:, H:]) * torch.tanh(g[:, 2 * H:] + pol.b_in + r(rz[:, :H] * pol.b_hn))
    return run(pol.actor, h)


@pytest.mark.parametrize("layout", ["sw128", "rowpad"])
@pytest.mark.parametrize("rtype,acts", [("GRU", ["Tanh", "ReLU"]), ("LSTM", ["LeakyReLU", "ReLU"]),
                                        ("RNN", ["LogSigmoid", "Tanh", "ELU"])])
def test_fused_policy_kernel_matches_its_restatement(cuda_device, rtype, acts, layout):
    import torch
    from madrona_bots import SimManager
    from madrona_bots.policy import FusedPolicy, SpeciesPolicy
    from prims_binding import shim
    torch.manual_seed(11)
    net, config = _make(rtype, acts)
    with torch.no_grad():                           # livelier logits than the default initialisation gives
        for p in net.parameters():
            p.mul_(2.0)
    pol = SpeciesPolicy(config, net.state_dict(), "cuda:0")
    fused = FusedPolicy(pol, layout)
    sim = SimManager(0, 40, 3, 32)
    _random_steps(sim, 5, 2)
    A = sim.num_agents()
    for start, end in ((0, A), (3, 3 + 129), (A - 77, A), (10, 11)):
        before = sim.action_tensor(False).to_torch().clone()
        idx, lg = sim.policy_act(start, end, fused, seed=5, draw=8, want_index=True, want_logits=True)
        ref = _bf16_reference_logits(pol, sim.pack_observations(start, end))
        scale = float(ref.abs().max())
        assert float((lg - ref).abs().max()) <= 3e-2 * scale + 1e-3, (float((lg - ref).abs().max()), scale)
        assert float((lg - ref).abs().mean()) <= 3e-3 * scale + 1e-4
        # sampling: the kernel's own logits -> softmax weights -> the documented inverse-CDF draw
        l = lg.cpu().numpy().astype(np.float32)
        w = np.exp((l - l.max(axis=1, keepdims=True)).astype(np.float32)).astype(np.float32)
        want = _host_sample(shim(), w, 5, 8, _agent_ids(sim, start, end))
        assert int((idx.cpu().numpy() != want).sum()) <= max(1, (end - start) // 500)
        act = sim.action_tensor(False).to_torch()
        assert torch.equal(act[start:end], torch.nn.functional.one_hot(idx.long(), 6).to(torch.int32))
        assert torch.equal(act[:start], before[:start]) and torch.equal(act[end:], before[end:])
    # sample = 0 leaves the actions alone
    before = sim.action_tensor(False).to_torch().clone()
    sim.policy_act(0, A, fused, 1, 1, sample=False, want_logits=True)
    assert torch.equal(sim.action_tensor(False).to_torch(), before)


@pytest.mark.parametrize("layout", ["sw128", "rowpad"])
def test_fused_policy_rejects_malformed_programs(cuda_device, layout):
    import torch
    from madrona_bots import SimManager
    from madrona_bots.policy import FusedPolicy, SpeciesPolicy
    net, config = _make("RNN", ["Tanh"])
    fused = FusedPolicy(SpeciesPolicy(config, net.state_dict(), "cuda:0"), layout)
    sim = SimManager(0, 4, 0, 32)
    sim.step()
    fused.program.ops[fused.program.n_ops - 1].w_ofs = fused.blob.numel()        # beyond the blob
    with pytest.raises(RuntimeError, match="beyond the blob"):
        sim.policy_act(0, 8, fused, 0, 0)
    fused.program.n_ops = 1
    with pytest.raises(RuntimeError, match="2..12 ops"):
        sim.policy_act(0, 8, fused, 0, 0)
    fused.program.layout = 7
    with pytest.raises(RuntimeError, match="unknown blob layout"):
        sim.policy_act(0, 8, fused, 0, 0)


@pytest.mark.parametrize("layout", ["sw128", "rowpad"])
def test_fused_policy_in_the_loop_matches_oracle(cuda_device, layout):
    import torch
    from madrona_bots.policy import PolicyRunner, SpeciesPolicy
    torch.manual_seed(1)
    pols = []
    for rtype, acts in (("GRU", ["Tanh", "ReLU"]), ("LSTM", ["LeakyReLU", "ReLU"]), ("RNN", ["LogSigmoid", "Tanh", "ReLU"]),
                        ("RNN", ["LogSigmoid", "Tanh", "LeakyReLU"])):
        net, config = _make(rtype, acts)
        pols.append(SpeciesPolicy(config, net.state_dict(), "cuda:0"))
    o, g = make_pair(32, 8, 32)
    o.step(); g.step()
    runner = PolicyRunner(g, pols, seed=4, fused=layout)
    for t in range(10):
        runner.act()
        act = g.action_tensor(False).to_torch()
        assert bool((act.sum(dim=1) == 1).all())
        o.tensor("ACTION")[:] = act.cpu().numpy()
        o.step(); g.step()
        compare(o, g, f"fused policy step {t}")
        o.shift_observations(); g.shift_observations()


@pytest.mark.parametrize("layout", ["sw128", "rowpad"])
def test_fused_policy_many_tiles_per_cta(cuda_device, layout):
    """~200 K rows in one call: every persistent CTA walks several tile pairs, so the weight ring, the barrier
    phases and the accumulator hand-over wrap around many times (the small cases above stay within one pass)."""
    import torch
    from madrona_bots import SimManager
    from madrona_bots.policy import FusedPolicy, SpeciesPolicy
    from prims_binding import shim
    torch.manual_seed(3)
    net, config = _make("GRU", ["Tanh", "ReLU"])
    with torch.no_grad():
        for p in net.parameters():
            p.mul_(2.0)
    pol = SpeciesPolicy(config, net.state_dict(), "cuda:0")
    fused = FusedPolicy(pol, layout)
    sim = SimManager(0, 6144, 9, 32)
    _random_steps(sim, 3, 4)
    A = sim.num_agents()
    assert A > 148 * 256 * 4
    idx, lg = sim.policy_act(0, A, fused, seed=2, draw=3, want_index=True, want_logits=True)
    ref = _bf16_reference_logits(pol, sim.pack_observations(0, A))
    scale = float(ref.abs().max())
    err = (lg - ref).abs()
    assert float(err.max()) <= 3e-2 * scale + 1e-3, (float(err.max()), scale, int(err.argmax()) // 6)
    assert float(err.mean()) <= 3e-3 * scale + 1e-4
    l = lg.cpu().numpy().astype(np.float32)
    w = np.exp((l - l.max(axis=1, keepdims=True)).astype(np.float32)).astype(np.float32)
    got = idx.cpu().numpy()
    for lo, hi in ((0, 20000), (A - 20000, A)):               # the host restatement is a Python loop: two slices
        want = _host_sample(shim(), w[lo:hi], 2, 3, _agent_ids(sim, lo, hi))
        assert int((got[lo:hi] != want).sum()) <= (hi - lo) // 500
    act = sim.action_tensor(False).to_torch()
    assert torch.equal(act, torch.nn.functional.one_hot(idx.long(), 6).to(torch.int32))
    # deterministic: the same call again gives the same picks and logits
    idx2, lg2 = sim.policy_act(0, A, fused, seed=2, draw=3, want_index=True, want_logits=True)
    assert torch.equal(idx, idx2) and torch.equal(lg, lg2)


def test_train_epoch_on_the_gpu_simulator(cuda_device):
    """the reference's loop body (learn/training_loop.py:36-136) runs end to end on the CUDA
    SimManager with the memory-input network the loop expects."""
    import torch
    from madrona_bots import SimManager
    from madrona_bots.learn import MemoryA2C, train_epoch
    torch.manual_seed(2)
    sim = SimManager(0, 64, 5, 32)
    nets = [MemoryA2C(rtype=t).to("cuda:0") for t in ("GRU", "LSTM", "RNN", "GRU")]
    opts = [torch.optim.Adam(n.parameters(), lr=1e-3) for n in nets]
    g = torch.Generator(device="cuda:0").manual_seed(3)
    for epoch in range(3):
        res = train_epoch(sim, nets, opts, generator=g)
        A = sim.num_agents()
        assert sum(r[2] for r in res) == A and all(np.isfinite(r[0]) and np.isfinite(r[1]) for r in res)
        act = sim.action_tensor(False).to_torch()
        assert act.shape == (A, 6) and bool((act.sum(dim=1) == 1).all())
        mem = sim.hidden_state_tensor(False).to_torch()
        assert float(mem.abs().max()) <= 1.0 and float(mem.abs().max()) > 0.0
        assert torch.equal(sim.position_tensor(True).to_torch(), sim.position_tensor(False).to_torch())


@pytest.mark.parametrize("layout", ["sw128", "rowpad"])
def test_fused_policy_matches_reference_models_py_on_shipped_weights(cuda_device, layout):
    """The fused acting kernel against the reference's OWN Python (learn/models.py ActorCritic.forward on the shipped
    checkpoints/universe_violence weights; fixture tests/golden/policy_universe_violence.npz, generator beside it):
    the 512 fixture observation rows are written into the simulator's exports, `mbots_policy_act` runs on them, and the
    logits must agree with the reference's fp32 logits to the stated bf16 tolerance: max |diff| <= 2e-2 of the logit
    scale, mean |diff| <= 4e-3 of it; the arg-max action agrees wherever the reference's top-2 gap exceeds twice the error."""
    import os
    import torch
    from madrona_bots import SimManager
    from madrona_bots.policy import FusedPolicy, load_universe_fixture
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "policy_universe_violence.npz")
    z = np.load(path)
    n = z["depth"].shape[0]
    sim = SimManager(0, 24, 1, 32)
    sim.step()
    assert sim.num_agents() >= n
    for name, key in (("SENSOR_DEPTH", "depth"), ("HEALTH", "health"), ("POSITION", "position"),
                      ("SENSOR_SEMANTIC", "semantic"), ("SURROUNDING", "surrounding")):
        sim.export(name).to_torch()[:n].copy_(torch.from_numpy(z[key]).cuda())
    want_obs = np.concatenate([z["depth"].astype(np.float32), z["health"], z["position"],
                               z["semantic"].astype(np.float32), z["surrounding"]], axis=1)
    np.testing.assert_array_equal(sim.pack_observations(0, n).cpu().numpy(), want_obs)
    for k, pol in enumerate(load_universe_fixture(path, "cuda:0"), 1):
        fused = FusedPolicy(pol, layout)
        lg = sim.policy_act(0, n, fused, seed=1, draw=k, sample=False, want_logits=True).cpu().numpy()
        ref = z[f"sp{k}.logits"]
        scale = float(np.abs(ref).max())
        err = np.abs(lg - ref)
        assert err.max() <= 2e-2 * scale, (k, layout, float(err.max()), scale)
        assert err.mean() <= 4e-3 * scale, (k, layout, float(err.mean()), scale)
        top2 = np.sort(ref, axis=1)[:, -2:]
        clear = (top2[:, 1] - top2[:, 0]) > 2.0 * err.max()       # the preferred action cannot flip where the gap exceeds the error
        assert (lg.argmax(1)[clear] == ref.argmax(1)[clear]).all()


def test_sampled_actions_do_not_depend_on_the_sharding(cuda_device):
    """an agent's draw is keyed on (seed, draw, GLOBAL world id, slot), not on its shard-local row, so worlds
    [8, 16) simulated alone sample the same actions -- and evolve identically with the policy in the loop -- as inside a
    16-world handle."""
    import os
    import torch
    from madrona_bots import SimManager
    from madrona_bots.policy import PolicyRunner, load_universe_fixture
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "policy_universe_violence.npz")
    pols = load_universe_fixture(path, "cuda:0")
    full = SimManager(0, 16, 21, 32)
    part = SimManager(0, 8, 21, 32, world_id_offset=8)
    runners = [PolicyRunner(full, pols, seed=3, fused="sw128"), PolicyRunner(part, pols, seed=3, fused="sw128")]
    for t in range(25):
        for sim, run in zip((full, part), runners):
            sim.step()
            run.act()
            sim.shift_observations()
    for sim in (full, part):
        sim.step()
    off = int(gpu_np(full, "DEBUG_WORLD_OFFSET")[8, 0])
    np.testing.assert_array_equal(gpu_np(full, "DEBUG_WORLD_COUNT")[8:], gpu_np(part, "DEBUG_WORLD_COUNT"))
    np.testing.assert_array_equal(gpu_np(full, "DEBUG_POSE")[off:], gpu_np(part, "DEBUG_POSE"))
    np.testing.assert_array_equal(gpu_np(full, "DEBUG_AGENT_HEALTH")[off:], gpu_np(part, "DEBUG_AGENT_HEALTH"))
    st = part.episode_stats()
    assert st[10] > 0 or st[11] > 0 or st[15] > 0, "the policy must have done something"


@pytest.mark.parametrize("one_launch", [True, False])
def test_act_async_rollout_equals_the_blocking_loop(cuda_device, one_launch):
    """mbots_policy_act_all (one launch for the four species) / mbots_policy_act_species (one per species): the acting
    kernels take their row ranges from the device, so act + step + sensor + shift
    can be enqueued steps ahead with no host wait. Same draws, same rows: the rollout must equal the blocking loop
    (step, act through species_blocks(), shift) on every export, and the indices / logits it reports (by observation row)
    must be the blocking call's."""
    import os
    import torch
    from madrona_bots import SimManager
    from madrona_bots.policy import PolicyRunner, load_universe_fixture
    from parity_util import ROW_EXPORTS, WORLD_EXPORTS, AGENT_DEBUG
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "policy_universe_violence.npz")
    pols = load_universe_fixture(path, "cuda:0")
    a, b = SimManager(0, 300, 5, 32), SimManager(0, 300, 5, 32)
    ra, rb = PolicyRunner(a, pols, seed=9, fused="sw128"), PolicyRunner(b, pols, seed=9, fused="sw128")
    cap = 300 * 64
    idx = torch.full((cap,), -1, dtype=torch.int32, device="cuda:0")
    lg = torch.zeros((cap, 6), dtype=torch.float32, device="cuda:0")
    a.step(); b.step()
    a.shift_observations(); b.shift_observations()
    steps = 24
    for t in range(steps):
        picked = ra.act(want_index=True)            # blocking cadence: the host knows the row ranges
        a.step()
        a.shift_observations()
        last = t == steps - 1
        rb.act_async(out_index=idx if last else None, out_logits=lg if last else None, one_launch=one_launch)   # no host wait
        b.rollout_step_async(None, shift=True)
        if last:
            b.sync()
            n_before = sum(int(p.numel()) for p in picked if p is not None)
            want = torch.cat([p for p in picked if p is not None]).cpu().numpy()
            np.testing.assert_array_equal(idx[:n_before].cpu().numpy(), want)
            assert bool((idx[n_before:] == -1).all()), "wrote index rows beyond the population"
            assert bool(lg[:n_before].abs().sum(dim=1).gt(0).all()) and not bool(lg[n_before:].any())
    b.sync()
    assert a.num_agents() == b.num_agents()
    for name in WORLD_EXPORTS + AGENT_DEBUG + ROW_EXPORTS + ["DEBUG_FOOD"]:
        np.testing.assert_array_equal(gpu_np(a, name), gpu_np(b, name), err_msg=name)
    st = b.episode_stats()
    assert st == a.episode_stats() and (st[10] > 0 or st[11] > 0 or st[15] > 0), "the policy must have done something"
    with pytest.raises(RuntimeError, match="SW128"):
        from madrona_bots.policy import FusedPolicy
        b.policy_act_species(0, FusedPolicy(pols[0], "rowpad"), 0, 0)
    with pytest.raises(RuntimeError, match="species must be"):
        b.policy_act_species(4, rb.fused[0], 0, 0)
