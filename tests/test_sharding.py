"""Multi-GPU host logic on the CPU (gloo, world_size 2): worlds shard as contiguous global-id ranges with no
data-path collective; the only exchange is the sum all-reduce of the 16-scalar episode-statistics vector.
The per-rank simulator here is the CPU oracle (same surface, same global-id keyed RNG streams as the CUDA
library), so the test proves: 2 shards + all-reduce == 1 unsharded run."""
import os
import socket
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (os.path.join(ROOT, "madrona-bots_b200"), HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

from madrona_bots.sharding import STAT_NAMES, allreduce_max, allreduce_stats, make_shard, shard_range  # noqa: E402

TOTAL_WORLDS, SEED, INIT, STEPS = 13, 17, 32, 25


def test_shard_range_partitions_exactly():
    for total in (1, 7, 8, 13, 65536):
        for k in (1, 2, 3, 4, 8):
            ranges = [shard_range(total, r, k) for r in range(k)]
            assert ranges[0][0] == 0 and ranges[-1][1] == total
            for (b0, e0), (b1, e1) in zip(ranges, ranges[1:]):
                assert e0 == b1 and e0 >= b0
            sizes = [e - b for b, e in ranges]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(8, 2, 2)
    assert len(STAT_NAMES) == 16


def test_allreduce_is_identity_without_a_process_group():
    assert allreduce_stats(list(range(16))) == list(range(16))
    assert allreduce_max(3.5) == 3.5


def _policy(sim, t):
    """Actions that depend only on an agent's own position, so sharded and unsharded runs agree on them."""
    pos = sim.tensor("POSITION")
    idx = ((pos[:, 0] * 5.0 + pos[:, 1] * 11.0).astype(np.int64) + t) % 6
    a = np.zeros((sim.num_agents(), 6), dtype=np.int32)
    a[np.arange(len(idx)), idx] = 1
    sim.tensor("ACTION")[:] = a


def _oracle_cls():
    from oracle_binding import OracleSim

    class Sim(OracleSim):                       # SimManager's constructor signature
        def __init__(self, gpu_id, num_worlds, rand_seed, init_agents, world_id_offset=0, **kw):
            super().__init__(num_worlds, rand_seed, init_agents, world_id_offset=world_id_offset)
    return Sim


def _run(sim):
    for t in range(STEPS):
        _policy(sim, t)
        sim.step()
        sim.shift_observations()
    return [int(x) for x in sim.episode_stats()]


def _worker(rank, world_size, port, out_dir):
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world_size)
    try:
        sim, (b, e) = make_shard(_oracle_cls(), rank, TOTAL_WORLDS, SEED, INIT, rank, world_size)
        assert sim.num_worlds == e - b
        local = _run(sim)
        total = allreduce_stats(local)
        slowest = allreduce_max(float(rank + 1))
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), local=np.array(local), total=np.array(total),
                 slowest=slowest, counts=np.array(sim.tensor("SPECIES_COUNT")), begin=b, end=e,
                 pose=np.array(sim.tensor("DEBUG_POSE")))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_two_shards_equal_one_run(tmp_path):
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "rank0.npz"), np.load(tmp_path / "rank1.npz")
    full = _oracle_cls()(0, TOTAL_WORLDS, SEED, INIT)
    want = _run(full)
    # every rank holds the same reduced vector, and it equals the unsharded run
    np.testing.assert_array_equal(r0["total"], r1["total"])
    np.testing.assert_array_equal(r0["total"], np.array(want))
    assert want[0] > 0 and want[1] == STEPS
    # additive entries really are sums of the two shards
    s = r0["local"] + r1["local"]
    s[1] = STEPS
    np.testing.assert_array_equal(s, np.array(want))
    assert float(r0["slowest"]) == 2.0 and float(r1["slowest"]) == 2.0
    # the shards tile the world range, and per-world results are those of the unsharded run
    assert (int(r0["begin"]), int(r0["end"]), int(r1["begin"]), int(r1["end"])) == (0, 7, 7, 13)
    np.testing.assert_array_equal(np.concatenate([r0["counts"], r1["counts"]]), full.tensor("SPECIES_COUNT"))
    np.testing.assert_array_equal(np.concatenate([r0["pose"], r1["pose"]]), full.tensor("DEBUG_POSE"))
