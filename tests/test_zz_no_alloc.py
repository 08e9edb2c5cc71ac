"""mbots_row_scan_static (include/mbots.h): the allocation-free variant of the step's row scan, called through the
C-ABI, must give exactly what the existing entry points give for the same populations -- mbots_tensor's
DEBUG_ROW_BASE / DEBUG_WORLD_OFFSET, mbots_species_blocks, mbots_num_agents of a handle that stepped them -- on calls
in a row that share the library's static scratch (sizes differ between calls, so stale scratch would show).
Integer work: bit-exact. This file sorts last on purpose."""
import ctypes

import numpy as np
import pytest

from parity_util import gpu_np

pytestmark = pytest.mark.gpu

MAX_WORLDS = 262144          # MBOTS_ROW_SCAN_STATIC_MAX_WORLDS


def _stepped_sim(num_worlds, seed, steps):
    """A handle whose worlds have diverged (births, deaths, respawns), driven through the existing entry points."""
    import torch
    from madrona_bots import SimManager
    sim = SimManager(0, num_worlds, seed, 24)
    rng = np.random.default_rng(seed)
    for _ in range(steps):
        acts = rng.integers(0, 2, (sim.num_agents(), 6)).astype(np.int32)     # any mix of the six action bits
        sim.action_tensor(False).to_torch().copy_(torch.from_numpy(acts))
        sim.step()
        sim.shift_observations()
    return sim


def _species_rows(sim):
    """(num_worlds, 4) int32 rows per (world, species), from the agent-order exports of the handle."""
    counts = gpu_np(sim, "DEBUG_WORLD_COUNT").reshape(-1).astype(np.int64)
    species = gpu_np(sim, "DEBUG_AGENT_SPECIES").reshape(-1).astype(np.int64)
    assert species.min() >= 1 and species.max() <= 4 and species.size == counts.sum()
    world = np.repeat(np.arange(counts.size), counts)
    rows = np.zeros((counts.size, 4), np.int32)
    np.add.at(rows, (world, species - 1), 1)
    return rows


def _expected(sim):
    blocks = [b for b, _ in sim.species_blocks()] + [sim.species_blocks()[-1][1]]
    totals = np.array(list(np.diff(blocks)) + [sim.num_agents(), gpu_np(sim, "DEBUG_WORLD_COUNT").max()], np.uint32)
    return gpu_np(sim, "DEBUG_ROW_BASE"), gpu_np(sim, "DEBUG_WORLD_OFFSET").reshape(-1), totals


def _call(lib, rows_np, dev):
    """One call of the new entry point on caller-owned device buffers; returns host copies of the three outputs."""
    import torch
    W = rows_np.shape[0]
    rows = torch.from_numpy(rows_np).to(dev)
    base = torch.full((W, 4), -7, dtype=torch.int32, device=dev)
    offs = torch.full((W,), -7, dtype=torch.int32, device=dev)
    tot = torch.full((6,), 7, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    rc = lib.mbots_row_scan_static(rows.data_ptr(), W, base.data_ptr(), offs.data_ptr(), tot.data_ptr(), stream)
    assert rc == 0, lib.mbots_last_error()
    torch.cuda.synchronize(dev)
    return base.cpu().numpy(), offs.cpu().numpy(), tot.cpu().numpy().view(np.uint32)


def test_static_row_scan_matches_the_handle_on_calls_in_a_row(cuda_device):
    import torch
    from madrona_bots import _native
    lib = _native.lib()
    lib.mbots_last_error.restype = ctypes.c_char_p
    big = _stepped_sim(2500, 11, 25)        # three tiles, the last one partial
    small = _stepped_sim(700, 12, 40)       # one partial tile: must not see the big call's tile records
    cases = [(big, _species_rows(big)), (small, _species_rows(small)), (big, _species_rows(big))]
    assert cases[0][1].sum(axis=1).min() != cases[0][1].sum(axis=1).max(), "worlds did not diverge"
    for n, (sim, rows) in enumerate(cases):
        want_base, want_offs, want_tot = _expected(sim)
        got_base, got_offs, got_tot = _call(lib, rows, cuda_device)
        np.testing.assert_array_equal(got_base, want_base, err_msg=f"call {n}: row_base")
        np.testing.assert_array_equal(got_offs, want_offs, err_msg=f"call {n}: world_offset")
        np.testing.assert_array_equal(got_tot, want_tot, err_msg=f"call {n}: totals")

    # back to back on one stream with no host wait in between, optional outputs left out
    rows_b, rows_s = (torch.from_numpy(c[1]).to(cuda_device) for c in cases[:2])
    out_b, out_s = torch.empty_like(rows_b), torch.empty_like(rows_s)
    stream = torch.cuda.current_stream(cuda_device).cuda_stream
    torch.cuda.synchronize(cuda_device)
    free_before = torch.cuda.mem_get_info(cuda_device)[0]
    for rows, out in ((rows_b, out_b), (rows_s, out_s), (rows_b, out_b)):
        assert lib.mbots_row_scan_static(rows.data_ptr(), rows.shape[0], out.data_ptr(), None, None, stream) == 0
    torch.cuda.synchronize(cuda_device)
    assert torch.cuda.mem_get_info(cuda_device)[0] == free_before, "the call changed the GPU's free memory"
    np.testing.assert_array_equal(out_b.cpu().numpy(), _expected(big)[0])
    np.testing.assert_array_equal(out_s.cpu().numpy(), _expected(small)[0])


def test_static_row_scan_rejects_what_its_scratch_cannot_hold(cuda_device):
    import torch
    from madrona_bots import _native
    lib = _native.lib()
    lib.mbots_last_error.restype = ctypes.c_char_p
    rows = torch.ones((MAX_WORLDS + 1024, 4), dtype=torch.int32, device=cuda_device)
    out = torch.zeros_like(rows)
    tot = torch.zeros(6, dtype=torch.int32, device=cuda_device)
    stream = torch.cuda.current_stream(cuda_device).cuda_stream
    assert lib.mbots_row_scan_static(rows.data_ptr(), MAX_WORLDS + 1, out.data_ptr(), None, None, stream) == -1
    assert b"MBOTS_ROW_SCAN_STATIC_MAX_WORLDS" in lib.mbots_last_error()
    assert lib.mbots_row_scan_static(None, 8, out.data_ptr(), None, None, stream) == -1
    assert lib.mbots_row_scan_static(rows.data_ptr(), 0, out.data_ptr(), None, None, stream) == -1
    assert lib.mbots_row_scan_static(rows.data_ptr() + 4, 8, out.data_ptr(), None, None, stream) == -1   # misaligned
    host = np.ones((8, 4), np.int32)
    assert lib.mbots_row_scan_static(host.ctypes.data, 8, out.data_ptr(), None, None, stream) == -1       # host memory
    # the largest size it accepts: every world holds one row of each species
    assert lib.mbots_row_scan_static(rows.data_ptr(), MAX_WORLDS, out.data_ptr(), None, tot.data_ptr(), stream) == 0
    torch.cuda.synchronize(cuda_device)
    w = np.arange(MAX_WORLDS, dtype=np.int64)[:, None]
    np.testing.assert_array_equal(out[:MAX_WORLDS].cpu().numpy(), (np.arange(4)[None, :] * MAX_WORLDS + w).astype(np.int32))
    np.testing.assert_array_equal(tot.cpu().numpy(), [MAX_WORLDS] * 4 + [4 * MAX_WORLDS, 4])
    assert not out[MAX_WORLDS:].any(), "wrote past num_worlds"
