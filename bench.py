#!/usr/bin/env python
"""bench.py -- agent-steps/s of the madrona-bots step hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = what learn/training_loop.py does to the simulator per epoch: write one action row per
live agent, SimManager.step() (Step + Sensor task graphs), SimManager.shift_observations().
Workload (config.workload): 65,536 independent worlds x 32 initial agents PER GPU (the configuration
BASELINE.json's metric is quoted on; weak scaling: every rank owns its own 65,536 global world ids),
uniform random one-hot actions from a pre-generated pool, seed 0.

  value        agent-steps/s, action pool resident in HBM (device-to-device write of the action rows)
  e2e          same loop driven through the C-ABI with HOST buffers: per step the action rows are
               copied from pinned host memory and the observation exports the caller reads
               (depth, semantic, position, health, surrounding, reward, species counts) are copied back;
               the copy-back of step t runs on a second stream under step t + 1 (two row sets, two host
               landing zones) and every copy, the last one included, is inside the timed region
  roofline     export+sensor kernel (the dominant one): algorithmic bytes / CUDA-event time
  cpu_baseline the CPU oracle (restated reference algorithm) on the host cores, bounded sample

--impl reference times the CPU oracle with all host threads on a bounded sample of the same workload
(the reference has no CPU backend and its GPU build does not compile here: SURVEY.md section 0).
Multi-GPU: launch with torchrun; one process per GPU, NCCL is used only for the statistics reduce.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (os.path.join(ROOT, "madrona-bots_b200"), os.path.join(ROOT, "tests"), ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)

METRIC = "agent-steps/sec"
UNIT = "agent-steps/s"
INIT_AGENTS = 32
SEED = 0
POOL = 4                       # pre-generated action buffers, cycled

# ALGORITHMIC bytes per agent-step (DESIGN.md "Algorithmic bytes"; SURVEY.md section 8(d) rows assigned
# to the kernel that owns them). STEP_BYTES = 321 in total.
ALGO_BYTES = {"step": 24 + 28 + 24 + 9.0, "export_sense": 44 + 64 + 64 + 64, "shift": 264}
# bytes the export kernel really has to move per agent given that rows move every step:
MOVED_BYTES_EXPORT = 280 + 280 + 36 + 64 + 4 + 2 + 41
# dram__bytes_read.sum + dram__bytes_write.sum of export_kernel<FAST> per agent, from the committed
# `ncu --set full` capture (profiles/r1k_ncu_full_summary.txt: 705.5 MB + 808.9 MB at 2,173,116 agents)
NCU_TRAFFIC_BYTES_PER_AGENT_EXPORT = (705.549568e6 + 808.883200e6) / 2173116.0
# same capture: what actually bounds the kernel is instruction issue
NCU_EXPORT_ISSUE = {"warp_instructions_per_world": 445786274 / 65536.0, "issue_active_frac": 0.765}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """SM clock, power and throttle reasons DURING the timed region: NVML polled every 5 ms from a thread
    (falls back to `nvidia-smi -lms 100` when pynvml is missing)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, gpu):
        self.gpu = gpu
        self.proc = None
        self.lines = []
        self.samples = []            # (sm_mhz, power_w, reasons bitmask)
        self.nvml = None
        self.stop_flag = False
        self.active = True           # samples are kept only while a timed region is running (pause / resume)

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.gpu)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        reasons_fn = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop_flag:
            if self.active:
                try:
                    smp = (float(n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)),
                           n.nvmlDeviceGetPowerUsage(self.h) / 1000.0, int(reasons_fn(self.h)))
                    if self.active:              # the region may have ended while NVML answered
                        self.samples.append(smp)
                except Exception:
                    pass
            time.sleep(0.005)

    def pause(self):
        self.active = False

    def resume(self):
        self.active = True

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.t.join(timeout=1)
            if not self.samples:
                return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["no samples"]}
            sm = sorted(x[0] for x in self.samples)
            mask = 0
            for x in self.samples:
                mask |= x[2]
            return {"sm_mhz": sm[len(sm) // 2], "sm_min_mhz": sm[0], "sm_max_mhz": self.max_mhz,
                    "power_w_max": max(x[1] for x in self.samples), "samples": len(sm), "source": "nvml, 5 ms, during both timed regions",
                    "reasons": sorted(k for k, b in self.BITS.items() if mask & b)}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(power),
                "samples": len(sm), "source": "nvidia-smi, 100 ms", "reasons": sorted(reasons)}


def one_hot_pool_numpy(rows, seed):
    import numpy as np
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(POOL):
        a = np.zeros((rows, 6), dtype=np.int32)
        a[np.arange(rows), rng.integers(0, 6, size=rows)] = 1
        out.append(a)
    return out


# ------------------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    """CPU oracle, all host threads, bounded sample of the workload."""
    if rank != 0:
        return
    import numpy as np
    from oracle_binding import OracleSim, oracle_lib
    lib = oracle_lib(False)
    lib.mbo_max_threads.restype = ctypes.c_int
    cores = int(lib.mbo_max_threads())
    sample_worlds = args.ref_worlds
    o = OracleSim(sample_worlds, SEED, INIT_AGENTS)
    pool = one_hot_pool_numpy(sample_worlds * 48, 1234)

    def step(i):
        A = o.num_agents()
        o.tensor("ACTION")[:] = pool[i % POOL][:A]
        o.step()
        o.shift_observations()
        return A

    for i in range(args.warmup):
        step(i)
    agent_steps = 0
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(i)
        agent_steps += o.num_agents()
    dt = time.perf_counter() - t0
    val = agent_steps / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32+i32",
        "data": "synthetic", "gpu_launches": 0,
        "config": workload_config(args, sample=f"{sample_worlds} of the {args.worlds_per_gpu} worlds per step"),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{sample_worlds} worlds x {INIT_AGENTS} agents, {args.steps} steps, "
                                   f"restated CPU oracle (reference has no CPU backend), {cores} threads"},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(args, sample=None):
    c = {"workload": f"{args.worlds_per_gpu} worlds x {INIT_AGENTS} initial agents per GPU, random one-hot actions, "
                     "step + sensor + shift_observations per step, one host synchronisation per step",
         "worlds_per_gpu": args.worlds_per_gpu, "init_agents_per_world": INIT_AGENTS, "seed": SEED,
         "sharding": f"worlds x{args.gpus} (independent worlds, no data-path collective)",
         "l2": "per-step working set (~1.5 GB at 65,536 worlds) exceeds the 126 MB L2; no explicit flush"}
    if sample:
        c["sample"] = sample
    return c


def cpu_baseline_sample(target_seconds=12.0):
    """The oracle on this box's host cores, bounded: ~10-20 s of CPU work."""
    import numpy as np
    from oracle_binding import OracleSim, oracle_lib
    lib = oracle_lib(False)
    lib.mbo_max_threads.restype = ctypes.c_int
    cores = int(lib.mbo_max_threads())
    worlds = 2048
    o = OracleSim(worlds, SEED, INIT_AGENTS)
    pool = one_hot_pool_numpy(worlds * 48, 1234)
    for i in range(3):
        o.tensor("ACTION")[:] = pool[i % POOL][:o.num_agents()]
        o.step(); o.shift_observations()
    agent_steps, steps = 0, 0
    t0 = time.perf_counter()
    while True:
        o.tensor("ACTION")[:] = pool[steps % POOL][:o.num_agents()]
        o.step(); o.shift_observations()
        agent_steps += o.num_agents()
        steps += 1
        dt = time.perf_counter() - t0
        if dt > target_seconds or steps >= 400:
            break
    return {"value": agent_steps / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{worlds} worlds x {INIT_AGENTS} agents, {steps} steps ({dt:.1f} s), restated CPU oracle, "
                      f"{cores} threads"}


def bind_host_to_gpu(local_rank):
    """Multi-GPU runs: pin this rank's host threads (and so its first-touch pinned staging buffers) to the
    CPUs NVML reports as local to its GPU, so that the e2e leg's PCIe copies do not cross the socket
    interconnect. Returns the number of CPUs bound to, or None when nothing was changed."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        ncpu = os.cpu_count() or 1
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        local = {i for i in range(ncpu) if (int(mask[i // 64]) >> (i % 64)) & 1}
        cpus = sorted(local & os.sched_getaffinity(0))
        if not cpus or len(cpus) == len(os.sched_getaffinity(0)):
            return None
        os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return None


# ------------------------------------------------------------------------------------------------
def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist
    from madrona_bots import SimManager
    from madrona_bots.sharding import allreduce_max, allreduce_stats

    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    bound_cpus = bind_host_to_gpu(local_rank) if world > 1 else None
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    W = args.worlds_per_gpu
    sim = SimManager(local_rank, W, SEED, INIT_AGENTS, world_id_offset=rank * W, max_agents_per_world=args.cap)
    rows_cap = W * 48                               # the population stays near 33 agents/world (max ~41)

    host_pool = [torch.from_numpy(a).pin_memory() for a in one_hot_pool_numpy(rows_cap, 1234 + rank)]
    dev_pool = [a.to(dev, non_blocking=True) for a in host_pool]
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def dev_step(i):
        A = sim.num_agents()
        assert A <= rows_cap, "action pool too small for the population"
        sim.action_tensor(False).to_torch().copy_(dev_pool[i % POOL][:A])
        sim.step_async()
        sim.shift_observations_async()
        sim.sync()                                    # one host synchronisation per step (the next step needs the row count)

    # ---------------- device-resident leg (value) ----------------
    for i in range(args.warmup):
        dev_step(i)
    sim.profile(True)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    st0 = sim.episode_stats()
    l0 = sim.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        dev_step(args.warmup + i)
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    if rank == 0:
        sampler.pause()                               # resumed for the end-to-end leg's timed region
    prof = sim.profile_read()
    sim.profile(False)
    st1 = sim.episode_stats()
    launches = sim.launch_count() - l0 - 1          # minus the stats-reduce kernel of st0/st1 bookkeeping
    local_agent_steps = st1[0] - st0[0]
    delta = [b - a for a, b in zip(st0, st1)]
    delta[1] = st1[1] - st0[1]
    tot = allreduce_stats(delta, device=dev)          # NCCL: the only collective of the path
    ms_max = allreduce_max(ms, device=dev)
    value = tot[0] / (ms_max * 1e-3)

    # ---------------- end-to-end leg (host buffers through the C-ABI) ----------------
    obs_names = [("SENSOR_DEPTH", 32), ("SENSOR_SEMANTIC", 32), ("POSITION", 8), ("HEALTH", 4), ("SURROUNDING", 8),
                 ("REWARD", 4)]
    # Two host landing zones: the device-to-host drain of step t runs on its own stream while step t + 1 is
    # computed (the exports of a step live in one of two row sets and are not written again before step t + 2,
    # which waits for the drain). The caller of step t + 1 finds the observations of step t on the host.
    host_out = [{n: torch.empty(rows_cap * b, dtype=torch.uint8).pin_memory() for n, b in obs_names} for _ in range(2)]
    host_sc = [torch.empty(W * 16, dtype=torch.uint8).pin_memory() for _ in range(2)]
    n_copy = max(1, int(os.environ.get("MBOTS_E2E_COPY_STREAMS", "1")))
    copy_streams = [torch.cuda.Stream(device=dev) for _ in range(n_copy)]
    drained = [None, None]                            # events: the drain that read row set (step % 2) is complete
    h2d = d2h = 0

    def e2e_step(i):
        nonlocal h2d, d2h
        main = torch.cuda.current_stream()
        A = sim.num_agents()
        sim.copy_from_host("ACTION", host_pool[i % POOL].data_ptr(), A * 24)
        h2d += A * 24
        if drained[i % 2] is not None:
            for ev in drained[i % 2]:
                main.wait_event(ev)                   # this step's export overwrites the row set drained two steps ago
        sim.step()
        sim.shift_observations_async()
        done = torch.cuda.Event()
        done.record(main)
        A = sim.num_agents()
        if drained[(i + 1) % 2] is not None:
            for ev in drained[(i + 1) % 2]:
                ev.synchronize()                      # the caller reads the observations of the previous step on the host
        for cs in copy_streams:
            cs.wait_event(done)
        for k, (n, b) in enumerate(obs_names):
            sim.copy_to_host(n, host_out[i % 2][n].data_ptr(), A * b, stream=copy_streams[k % n_copy])
            d2h += A * b
        sim.copy_to_host("SPECIES_COUNT", host_sc[i % 2].data_ptr(), W * 16, stream=copy_streams[0])
        d2h += W * 16
        drained[i % 2] = []
        for cs in copy_streams:
            ev = torch.cuda.Event()
            ev.record(cs)
            drained[i % 2].append(ev)

    for i in range(min(args.warmup, 3)):
        e2e_step(i)
    h2d = d2h = 0
    barrier()
    se0 = sim.episode_stats()
    if rank == 0:
        sampler.resume()
    e0.record()
    for i in range(args.steps):
        e2e_step(min(args.warmup, 3) + i)
    for cs in copy_streams:
        torch.cuda.current_stream().wait_stream(cs)           # the last drain is inside the timed region
    e1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None    # sampled during both timed regions (device-resident and end-to-end)
    ms_e = allreduce_max(e0.elapsed_time(e1), device=dev)
    se1 = sim.episode_stats()
    tot_e = allreduce_stats([b - a for a, b in zip(se0, se1)], device=dev)
    e2e_value = tot_e[0] / (ms_e * 1e-3)

    if rank != 0:
        return None
    peak, peak_src = measured_peaks()
    agents_per_launch = local_agent_steps / max(1, args.steps)
    k_ms = prof["export_sense"]["ms"]
    achieved = ALGO_BYTES["export_sense"] * agents_per_launch / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0
    kernels = {}
    for name in ("step", "scan", "export_sense", "shift"):
        kms = prof[name]["ms"]
        ent = {"ms": kms, "calls": prof[name]["calls"]}
        if name in ALGO_BYTES and kms > 0:
            ent["algo_gbs"] = ALGO_BYTES[name] * agents_per_launch / (kms * 1e-3) / 1e9
            ent["frac"] = ent["algo_gbs"] / peak
        kernels[name] = ent
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f32+i32", "data": "synthetic",
        "config": dict(workload_config(args), host_affinity=(f"rank bound to its GPU's {bound_cpus} local CPUs (NVML)"
                                                              if bound_cpus else "unchanged")),
        "clocks": clocks, "gpu_launches": int(launches),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d // max(1, args.steps),
                "d2h_bytes_per_step": d2h // max(1, args.steps), "ms_per_step": ms_e / args.steps,
                "overlap": "device-to-host drain of step t on a copy stream under step t+1; all copies inside the timed region"},
        "roofline": {"bound": "hbm", "kernel": "export_kernel (row carry + observation export + 32-ray sensor)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": NCU_TRAFFIC_BYTES_PER_AGENT_EXPORT * agents_per_launch,
                     "traffic_source": "ncu --set full capture under profiles/ (bytes per agent x agents of this run)",
                     "peak_source": peak_src,
                     "algo_bytes_per_agent": ALGO_BYTES["export_sense"],
                     "moved_bytes_per_agent": MOVED_BYTES_EXPORT,
                     "achieved_moved": MOVED_BYTES_EXPORT * agents_per_launch / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0,
                     "agents_per_launch": agents_per_launch, "kernel_ms": k_ms,
                     "note": "issue-bound, not HBM-bound (32-ray sensor arithmetic): see issue_*",
                     "issue_active_frac_ncu": NCU_EXPORT_ISSUE["issue_active_frac"],
                     "issue_warp_instructions_per_world_ncu": NCU_EXPORT_ISSUE["warp_instructions_per_world"]},
        "kernels": kernels,
        "episode_stats": dict(zip(("agent_steps", "steps", "births", "deaths", "respawns", "overflow"),
                                  (tot[0], tot[1], tot[10], tot[11], tot[12], tot[13]))),
    }
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_sample()
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--worlds", type=int, default=65536,
                    help="TOTAL worlds of the job, sharded over the GPUs (BASELINE.json configs[3]: strong scaling)")
    ap.add_argument("--worlds-per-gpu", type=int, default=0,
                    help="fixed worlds per GPU instead (weak scaling); 0 = --worlds // --gpus")
    ap.add_argument("--ref-worlds", type=int, default=4096, help="worlds per step of the CPU reference arm")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cap", type=int, default=0, help="per-world slot capacity (0 = library default, 128)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    args.scaling = "weak" if args.worlds_per_gpu > 0 else "strong"
    if args.worlds_per_gpu <= 0:
        args.worlds_per_gpu = max(1, args.worlds // max(1, args.gpus))

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    # stdout carries exactly one JSON line: while the run lasts, file descriptor 1 points at stderr, so that
    # banners written by native libraries (NCCL prints its version to stdout under NCCL_DEBUG) land there
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    line = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        line = run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
    if line is not None:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
