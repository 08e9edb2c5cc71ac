#!/usr/bin/env python
"""bench.py -- agent-steps/s of the madrona-bots step hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = what learn/training_loop.py does to the simulator per epoch: write one action row per
live agent, SimManager.step() (Step + Sensor task graphs), SimManager.shift_observations().
Workload (config.workload): 65,536 independent worlds x 32 initial agents IN TOTAL, sharded over the GPUs
(BASELINE.json metric / configs[3]: strong scaling, 65,536 // N worlds per GPU, rank r owns global world ids
[r * W/N, (r+1) * W/N)); uniform random one-hot actions from a pre-generated pool, seed 0. Multi-GPU runs add a
companion leg with 65,536 worlds on EVERY GPU under the key weak_65536_per_gpu.

  value        agent-steps/s, action pool resident in HBM: one CUDA-graph launch per step (action write with the
               row count resolved on the device, step + sensor, shift), the host up to 3 steps ahead of the device
  value_sync_per_step  same work with the reference's blocking cadence (the host waits for every step)
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
# the fused rollout step: the export kernel also leaves prev := cur (the shift call's 132 B read + 132 B written per agent)
ALGO_BYTES["export_sense_fused_shift"] = ALGO_BYTES["export_sense"] + ALGO_BYTES["shift"]
# measured DRAM traffic and warp instructions come from profiles/ncu_capture.json (ncu_capture() below), never from constants

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
    c = {"workload": f"{args.worlds_per_gpu} worlds x {INIT_AGENTS} initial agents per GPU "
                     f"({args.worlds_per_gpu * args.gpus} worlds in total, {args.scaling} scaling), random one-hot actions, "
                     "action write + step + sensor + shift_observations per step as one CUDA-graph launch, host up to 3 steps ahead",
         "worlds_per_gpu": args.worlds_per_gpu, "init_agents_per_world": INIT_AGENTS, "seed": SEED,
         "sharding": f"worlds x{args.gpus} (independent worlds, no data-path collective)",
         "l2": f"per-step working set (~{args.worlds_per_gpu * 24 // 1000} MB: two row sets + state) "
               + ("exceeds the 126 MB L2; no explicit flush" if args.worlds_per_gpu >= 8192 else
                  "fits the 126 MB L2 at this size: steps run warm, as they do in a rollout")}
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
def source_sha():
    """Hash of the kernel sources an ncu capture belongs to (profiles/ncu_capture.json is only trusted for the code it saw)."""
    import hashlib
    h = hashlib.sha256()
    for f in ("sim_kernels.cu", "sim_prims.cuh", "sim_dev.h"):
        with open(os.path.join(ROOT, "madrona-bots_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def ncu_capture():
    """profiles/ncu_capture.json (written by tools/ncu_capture_json.py from an `ncu --set full` report of this very code):
    per kernel the DRAM bytes and warp instructions of one launch. None when absent or taken from other sources."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_capture.json")) as f:
            cap = json.load(f)
    except Exception:
        return None, "no capture (profiles/ncu_capture.json missing)"
    if cap.get("source_sha") != source_sha():
        return None, f"STALE capture: taken from sources {cap.get('source_sha')}, this tree is {source_sha()} -- re-run tools/gpu_ncu_capture.sh"
    return cap, "profiles/ncu_capture.json (same sources)"


class Leg:
    """K-step timed regions, repeated until ~0.25 s of device time has been measured; the median repeat is reported."""

    def __init__(self, world, dev):
        self.world, self.dev = world, dev

    def barrier(self):
        import torch
        import torch.distributed as dist
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run(self, sim, step_fn, finish_fn, steps, first_index, max_repeats=20, min_seconds=0.25):
        import torch
        from madrona_bots.sharding import allreduce_max, allreduce_stats
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        out, idx, repeats = [], first_index, None
        while repeats is None or len(out) < repeats:
            self.barrier()
            st0 = sim.episode_stats()
            e0.record()
            for i in range(steps):
                step_fn(idx + i)
            finish_fn()
            e1.record()
            self.barrier()
            idx += steps
            st1 = sim.episode_stats()
            ms = allreduce_max(e0.elapsed_time(e1), device=self.dev)
            tot = allreduce_stats([b - a for a, b in zip(st0, st1)], device=self.dev)     # NCCL: the only collective of the path
            out.append({"ms": ms, "agent_steps": tot[0], "local_agent_steps": st1[0] - st0[0], "stats": tot})
            if repeats is None:
                repeats = int(min(max_repeats, max(1, -(-min_seconds * 1e3 // max(ms, 1e-3)))))
        rates = sorted(range(len(out)), key=lambda k: out[k]["agent_steps"] / out[k]["ms"])
        med = out[rates[len(rates) // 2]]
        med = dict(med, repeats=len(out), value=med["agent_steps"] / (med["ms"] * 1e-3),
                   spread=[out[rates[0]]["agent_steps"] / (out[rates[0]]["ms"] * 1e-3),
                           out[rates[-1]]["agent_steps"] / (out[rates[-1]]["ms"] * 1e-3)])
        return med, idx


def link_probe(dev, sizes, seconds=1.0):
    """Pinned device-to-host copies of the e2e leg's sizes, nothing else running: what the host link of this rank
    delivers while every rank probes at once (the ceiling of the e2e leg)."""
    import torch
    src = [torch.empty(n, dtype=torch.uint8, device=dev) for n in sizes]
    dst = [torch.empty(n, dtype=torch.uint8).pin_memory() for n in sizes]
    for a, b in zip(src, dst):
        b.copy_(a, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps, t0 = 0, time.perf_counter()
    e0.record()
    while time.perf_counter() - t0 < seconds:
        for a, b in zip(src, dst):
            b.copy_(a, non_blocking=True)
        torch.cuda.synchronize()
        reps += 1
    e1.record()
    torch.cuda.synchronize()
    return sum(sizes) * reps / (e0.elapsed_time(e1) * 1e-3) / 1e9


def policy_leg(args, sim, leg, dev, world, idx):
    """learn/training_loop.py:36-83,135-137 without the learner: step(), then per species ONE fused acting kernel
    (construct_obs -> species network on tcgen05 tensor cores -> softmax -> Categorical sample -> one-hot action rows,
    csrc/policy_umma.cu), then shift_observations(). Weights: the shipped checkpoints/universe_violence networks
    (fixture tests/golden/policy_universe_violence.npz, written by make_policy_golden.py from the reference's files)."""
    import torch
    from madrona_bots.policy import PolicyRunner, load_universe_fixture
    from madrona_bots.sharding import allreduce_max, allreduce_stats
    path = os.path.join(ROOT, "tests", "golden", "policy_universe_violence.npz")
    pols = load_universe_fixture(path, dev)
    runner = PolicyRunner(sim, pols, seed=SEED, fused="sw128")
    flops_per_agent = []
    for p in pols:                                    # useful multiply-adds of one agent's forward pass (actor path)
        macs = sum(l[1].shape[0] * l[1].shape[1] for l in p.feature if l[0] == "linear")
        macs += p.w_ih_t.shape[0] * (p.hidden_dim * (3 if p.rtype in ("GRU", "LSTM") else 1))   # LSTM: f gate unused from c0 = 0
        macs += sum(l[1].shape[0] * l[1].shape[1] for l in p.actor if l[0] == "linear")
        flops_per_agent.append(2.0 * macs)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    t_pol = t_all = 0.0
    flops = 0.0
    agent_steps = 0
    for t in range(args.warmup + args.steps):
        ev[0].record()
        sim.step()
        ev[1].record()
        blocks = sim.species_blocks()
        runner.act()
        ev[2].record()
        sim.shift_observations_async()
        ev[3].record()
        torch.cuda.synchronize()
        if t >= args.warmup:
            t_pol += ev[1].elapsed_time(ev[2])
            t_all += ev[0].elapsed_time(ev[3])
            agent_steps += sim.num_agents()
            flops += sum(f * (b - a) for f, (a, b) in zip(flops_per_agent, blocks))
    tot = allreduce_stats([agent_steps, 0, int(flops / 1e6)], device=dev)
    t_all_max = allreduce_max(t_all, device=dev)
    t_pol_max = allreduce_max(t_pol, device=dev)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            tf_peak, tf_src = float(json.load(f)["bf16_tflops_sustained"]), "MEASURED_PEAKS.json bf16_tflops_sustained"
    except Exception:
        tf_peak, tf_src = 1400.0, "fallback"
    tflops = tot[2] * 1e6 / (t_pol_max * 1e-3) / 1e12 / world          # per GPU

    # the same loop with nothing waited for: the acting kernels take their row ranges from the device
    # (mbots_policy_act_species), the step is the one-launch rollout step with the shift fused, the host runs ahead
    def run_async(n):
        inflight = []
        for _ in range(n):
            runner.act_async()
            sim.rollout_step_async(None, shift=True)
            e = torch.cuda.Event()
            e.record()
            inflight.append(e)
            if len(inflight) > RUN_AHEAD:
                inflight.pop(0).synchronize()
        sim.sync()

    run_async(args.warmup)
    a_ms, a_agents = [], 0
    for _ in range(5):
        st0 = sim.episode_stats()[0]
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        run_async(args.steps)
        e1.record()
        torch.cuda.synchronize()
        a_ms.append(allreduce_max(e0.elapsed_time(e1), device=dev))
        # agents acted on = the populations BEFORE each of the K steps (agent_steps counts them after)
        a_agents = sim.episode_stats()[0] - st0
    a_ms.sort()
    a_med = a_ms[len(a_ms) // 2]
    a_tot = allreduce_stats([a_agents, 0], device=dev)[0]
    return {"value": tot[0] / (t_all_max * 1e-3), "unit": UNIT, "ms_per_step": t_all_max / args.steps,
            "policy_ms_per_step": t_pol_max / args.steps, "steps": args.steps,
            "async": {"value": a_tot / (a_med * 1e-3), "unit": UNIT, "ms_per_step": a_med / args.steps,
                      "launches_per_step": 2, "run_ahead_steps": RUN_AHEAD,
                      "what": "act_async (ONE acting kernel for the four species, row ranges resolved on the device) + "
                              "rollout_step_async(shift) (one graph launch): the same work with no host wait, host up to 3 steps "
                              "ahead; median of 5 regions of K steps"},
            "what": "step() + 4 fused acting kernels (tcgen05, bf16 in / fp32 accumulate) + shift_observations per step; "
                    "the host waits for every step (the species row ranges come from it)",
            "weights": "checkpoints/universe_violence latest_model_epoch_788 (GRU, LSTM, RNN, RNN), fixture npz",
            "roofline": {"bound": "tensor", "kernel": "policy_umma_kernel (4 launches per step, one per species)",
                         "achieved": tflops, "peak": tf_peak, "unit": "TFLOP/s", "frac": tflops / tf_peak, "peak_source": tf_src,
                         "flops_per_agent": flops_per_agent,
                         "note": "useful FLOPs of the networks (no padding); the time includes the four launches and the "
                                 "host's species_blocks() read between them"}}


RUN_AHEAD = 3          # steps the host may have in flight in the device-resident leg


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
    leg = Leg(world, dev)

    host_pool = [torch.from_numpy(a).pin_memory() for a in one_hot_pool_numpy(rows_cap, 1234 + rank)]
    dev_pool = [a.to(dev, non_blocking=True) for a in host_pool]
    torch.cuda.synchronize()

    # ---------------- device-resident leg (value): one graph launch per step, host runs ahead ----------------
    inflight = []

    def dev_step(i):
        # action rows of the pool -> action export (live row count read on the device) -> step + sensor -> shift:
        # one CUDA graph launch; the host waits only when RUN_AHEAD steps are in flight
        sim.rollout_step_async(dev_pool[i % POOL], shift=True)
        ev = torch.cuda.Event()
        ev.record()
        inflight.append(ev)
        if len(inflight) > RUN_AHEAD:
            inflight.pop(0).synchronize()

    def dev_finish():
        sim.sync()
        inflight.clear()

    for i in range(args.warmup):
        dev_step(i)
    dev_finish()
    assert sim.num_agents() <= rows_cap, "action pool too small for the population"
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    l0, g0 = sim.launch_count(), sim.graph_launch_count()
    res, idx = leg.run(sim, dev_step, dev_finish, args.steps, args.warmup)
    launches_per_step = (sim.launch_count() - l0 - 2 * res["repeats"]) / (res["repeats"] * args.steps)   # minus the stats reads
    graph_launches = sim.graph_launch_count() - g0
    if rank == 0:
        sampler.pause()

    # the same steps with the reference's blocking call sequence: write, step() (waits), shift
    def sync_step(i):
        sim.write_actions(dev_pool[i % POOL])
        sim.step_async()
        sim.shift_observations_async()
        sim.sync()

    res_sync, idx = leg.run(sim, sync_step, lambda: None, args.steps, idx, max_repeats=5, min_seconds=0.1)

    # per-kernel device times: the timed loop's own kernels with events between them (plain launches instead of the
    # graph), then the unfused call sequence (separate shift pass) for comparison
    sim.profile(True)
    for i in range(args.steps):
        sim.rollout_step_async(dev_pool[(idx + i) % POOL], shift=True)
        sim.sync()
    idx += args.steps
    torch.cuda.synchronize()
    prof = sim.profile_read()
    sim.profile(True)                                 # resets the accumulators
    for i in range(args.steps):
        sync_step(idx + i)
    idx += args.steps
    torch.cuda.synchronize()
    prof_unfused = sim.profile_read()
    sim.profile(False)
    prof_agents = sim.num_agents()

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
    moved = {"h2d": 0, "d2h": 0}

    def e2e_step(i):
        main = torch.cuda.current_stream()
        A = sim.num_agents()
        sim.copy_from_host("ACTION", host_pool[i % POOL].data_ptr(), A * 24)
        moved["h2d"] += A * 24
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
            moved["d2h"] += A * b
        sim.copy_to_host("SPECIES_COUNT", host_sc[i % 2].data_ptr(), W * 16, stream=copy_streams[0])
        moved["d2h"] += W * 16
        drained[i % 2] = []
        for cs in copy_streams:
            ev = torch.cuda.Event()
            ev.record(cs)
            drained[i % 2].append(ev)

    def e2e_finish():
        for cs in copy_streams:
            torch.cuda.current_stream().wait_stream(cs)           # the last drain is inside the timed region

    for i in range(3):
        e2e_step(idx + i)
    idx += 3
    e2e_finish()
    moved["h2d"] = moved["d2h"] = 0
    if rank == 0:
        sampler.resume()
    res_e, idx = leg.run(sim, e2e_step, e2e_finish, args.steps, idx, max_repeats=5, min_seconds=0.2)
    clocks = sampler.stop() if rank == 0 else None    # sampled during the device-resident and the end-to-end timed regions
    e2e_steps = res_e["repeats"] * args.steps
    d2h_per_step = moved["d2h"] // max(1, e2e_steps)
    leg.barrier()
    link = link_probe(dev, [prof_agents * b for _, b in obs_names] + [W * 16], seconds=1.0)
    link_all = allreduce_stats([int(link * 1e6), 0], device=dev)[0] / 1e6 if world > 1 else link
    link_min = -allreduce_max(-link, device=dev) if world > 1 else link

    # ---------------- policy in the loop (BASELINE.json configs[4]: universe_violence policy, inference part) ----------------
    policy = None
    if args.policy:
        policy = policy_leg(args, sim, leg, dev, world, idx)

    # ---------------- weak-scaling companion: 65,536 worlds on every GPU ----------------
    weak = None
    if world > 1 and args.scaling == "strong" and not args.no_weak_leg:
        del sim
        torch.cuda.empty_cache()
        Ww = args.worlds
        simw = SimManager(local_rank, Ww, SEED, INIT_AGENTS, world_id_offset=rank * Ww, max_agents_per_world=args.cap)
        poolw = [torch.from_numpy(a).to(dev) for a in one_hot_pool_numpy(Ww * 48, 99 + rank)]
        pend = []

        def w_step(i):
            simw.rollout_step_async(poolw[i % POOL], shift=True)
            ev = torch.cuda.Event()
            ev.record()
            pend.append(ev)
            if len(pend) > RUN_AHEAD:
                pend.pop(0).synchronize()

        def w_finish():
            simw.sync()
            pend.clear()

        for i in range(args.warmup):
            w_step(i)
        w_finish()
        rw, _ = leg.run(simw, w_step, w_finish, args.steps, args.warmup, max_repeats=5, min_seconds=0.1)
        weak = {"value": rw["value"], "unit": UNIT, "ms_per_step": rw["ms"] / args.steps, "worlds_per_gpu": Ww,
                "scaling": "weak", "repeats": rw["repeats"]}

    if rank != 0:
        return None
    peak, peak_src = measured_peaks()
    agents_per_launch = res["local_agent_steps"] / max(1, args.steps)
    k_ms = prof["export_sense"]["ms"]
    algo_export = ALGO_BYTES["export_sense_fused_shift"]
    achieved = algo_export * prof_agents / (k_ms * 1e-3) / 1e9 if k_ms > 0 else 0.0

    def kernel_table(pr, names, algo):
        out = {}
        for name in names:
            kms = pr[name]["ms"]
            ent = {"ms": kms, "calls": pr[name]["calls"]}
            if name in algo and kms > 0:
                ent["algo_bytes_per_agent"] = algo[name]
                ent["algo_gbs"] = algo[name] * prof_agents / (kms * 1e-3) / 1e9
                ent["frac"] = ent["algo_gbs"] / peak
            out[name] = ent
        return out

    kernels = kernel_table(prof, ("step", "scan", "export_sense"), dict(ALGO_BYTES, export_sense=algo_export))
    kernels_unfused = kernel_table(prof_unfused, ("step", "scan", "export_sense", "shift"), ALGO_BYTES)
    cap, cap_src = ncu_capture()
    traffic = issue_frac = inst_per_world = None
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    if cap and "export_kernel" in cap.get("kernels", {}):
        ck = cap["kernels"]["export_kernel"]
        traffic = (ck["dram_read_bytes"] + ck["dram_write_bytes"]) / ck["agents"] * prof_agents
        inst_per_world = ck["warp_instructions"] / ck["worlds"]
        if k_ms > 0:
            issue_frac = inst_per_world * W / (k_ms * 1e-3) / (148 * 4 * sm_mhz * 1e6)
    line = {
        "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": res["ms"] / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f32+i32", "data": "synthetic",
        "config": dict(workload_config(args), host_affinity=(f"rank bound to its GPU's {bound_cpus} local CPUs (NVML)"
                                                              if bound_cpus else "unchanged")),
        "timing": {"repeats": res["repeats"], "reported": "median repeat of K steps", "spread": res["spread"],
                   "run_ahead_steps": RUN_AHEAD},
        "clocks": clocks, "gpu_launches": int(round(launches_per_step * args.steps)),
        "launches_per_step": launches_per_step, "graph_launches": int(graph_launches),
        "value_sync_per_step": {"value": res_sync["value"], "ms_per_step": res_sync["ms"] / args.steps,
                                "what": "same work, the host waits for every step (write_actions, step_async, shift, sync): "
                                        "the reference's blocking step() cadence"},
        "e2e": {"value": res_e["value"], "unit": UNIT, "h2d_bytes_per_step": moved["h2d"] // max(1, e2e_steps),
                "d2h_bytes_per_step": d2h_per_step, "ms_per_step": res_e["ms"] / args.steps,
                "overlap": "device-to-host drain of step t on a copy stream under step t+1; all copies inside the timed region",
                "link_gbs_measured": link, "link_gbs_all_ranks": link_all, "link_gbs_slowest_rank": link_min,
                "link_frac": (d2h_per_step / (res_e["ms"] / args.steps * 1e-3) / 1e9) / link if link > 0 else None,
                "link_probe": "1 s of pinned device-to-host copies of the same sizes, every rank at once, nothing else running"},
        "roofline": {"bound": "issue", "kernel": "export_kernel (row carry + observation export + 32-ray sensor + fused shift)",
                     "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "traffic_source": cap_src, "peak_source": peak_src,
                     "algo_bytes_per_agent": algo_export,
                     "algo_bytes_what": "236 B export share of STEP_BYTES (exports 44 W, sensors 64 W, prev-sensor carry 64 R + 64 W) "
                                        "+ 264 B SHIFT_BYTES: the shift call is fused into this kernel (SURVEY.md 8(d))",
                     "agents_per_launch": prof_agents, "kernel_ms": k_ms,
                     "kernel_ms_source": "CUDA events between the kernels on the launch stream, K steps right after the timed region",
                     "issue_frac": issue_frac, "issue_peak": "148 SMs x 4 schedulers x sampled SM clock (warp instructions / s)",
                     "warp_instructions_per_world": inst_per_world,
                     "note": "issue-bound (32-ray sensor arithmetic): frac is the algorithmic-HBM figure the contract asks for, "
                             "issue_frac the fraction of the instruction-issue roofline"},
        "kernels": kernels, "kernels_unfused_call_sequence": kernels_unfused,
        "episode_stats": dict(zip(("agent_steps", "steps", "births", "deaths", "respawns", "overflow"),
                                  (res["stats"][0], res["stats"][1], res["stats"][10], res["stats"][11], res["stats"][12],
                                   res["stats"][13]))),
    }
    if weak:
        line["weak_65536_per_gpu"] = weak
    if policy:
        line["policy_rollout"] = policy
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
    ap.add_argument("--no-weak-leg", action="store_true", help="multi-GPU: skip the 65,536-worlds-per-GPU companion leg")
    ap.add_argument("--policy", action="store_true",
                    help="add the policy-in-the-loop leg (universe_violence networks through the fused tcgen05 acting kernel)")
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
