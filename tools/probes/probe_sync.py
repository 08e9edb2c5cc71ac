"""Where does a blocking step spend its host time? (debug probe)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "madrona-bots_b200"))
import torch
from madrona_bots import SimManager
W = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
sim = SimManager(0, W, 0, 32)
pool = torch.zeros((W * 48, 6), dtype=torch.int32, device="cuda"); pool[:, 0] = 1
def t(fn, n=30):
    for _ in range(5): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
def a(): sim.step_async(); sim.sync()
def b(): sim.write_actions(pool); sim.step_async(); sim.sync()
def c(): sim.write_actions(pool); sim.step_async(); sim.shift_observations_async(); sim.sync()
def d(): sim.rollout_step_async(pool, shift=True); sim.sync()
def e(): sim.step_async(); sim.shift_observations_async(); sim.sync()
def f(): sim.step_async(); sim.shift_observations_async(); torch.cuda.synchronize()
for name, fn in (("step+sync", a), ("write+step+sync", b), ("write+step+shift+sync", c), ("fused+sync", d), ("step+shift+sync", e), ("step+shift+devsync", f)):
    print(W, name, round(t(fn), 4), "ms", flush=True)
