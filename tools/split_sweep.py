"""A/B of the export kernel's split settings: the bench's device-resident loop only (one graph launch per step, host
three steps ahead), one process per setting because the library reads its switches once.
    python tools/split_sweep.py WORLDS [STEPS]   -> one JSON line: ms per step, export kernel ms (events), population stats
Environment: MBOTS_EXPORT_SPLIT_LOG2 (csrc/sim_kernels.cu export_split_log2); MBOTS_EXPORT_SPLIT_HEAVY64 only existed in the
r3c experiment (profiles/r2c_export_lanes_negative_result.txt, fifth entry)."""
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "madrona-bots_b200"))


def main():
    import torch
    import bench
    from madrona_bots import SimManager
    W = int(sys.argv[1])
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    prewarm = int(os.environ.get("SWEEP_PREWARM", "10"))
    use_events = os.environ.get("SWEEP_EVENTS", "1") != "0"     # a torch event per step bounds the host's run-ahead
    dev = torch.device("cuda", 0)
    sim = SimManager(0, W, bench.SEED, bench.INIT_AGENTS)
    rows_cap = W * 48
    pool = [torch.from_numpy(a).to(dev) for a in bench.one_hot_pool_numpy(rows_cap, 1234)]
    idx = 0
    for _ in range(prewarm):
        sim.rollout_step_async(pool[idx % bench.POOL], shift=True)
        idx += 1
    sim.sync()
    times = []
    for _ in range(9):
        inflight = []
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            sim.rollout_step_async(pool[idx % bench.POOL], shift=True)
            idx += 1
            if use_events:
                ev = torch.cuda.Event()
                ev.record()
                inflight.append(ev)
                if len(inflight) > bench.RUN_AHEAD:
                    inflight.pop(0).synchronize()
        e1.record()
        sim.sync()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) / steps)
    sim.profile(True)
    for _ in range(steps):
        sim.rollout_step_async(pool[idx % bench.POOL], shift=True)
        sim.sync()
        idx += 1
    prof = sim.profile_read()
    sim.profile(False)
    counts = sim.export("DEBUG_WORLD_COUNT").to_torch().float()
    print(json.dumps({
        "worlds": W, "split_log2": os.environ.get("MBOTS_EXPORT_SPLIT_LOG2", "default"),
        "graphs": os.environ.get("MBOTS_GRAPHS", "1"), "events": use_events,
        "ms_per_step": round(statistics.median(times), 5), "ms_min": round(min(times), 5),
        "kernels_ms": {k: round(v["ms"], 5) for k, v in prof.items()},
        "agents": sim.num_agents(), "n_mean": round(float(counts.mean()), 2), "n_std": round(float(counts.std()), 2),
        "n_max": int(counts.max()), "steps_done": idx}))


if __name__ == "__main__":
    main()
