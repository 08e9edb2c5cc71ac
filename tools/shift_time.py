"""Device time of shift_observations() (CUDA events around the kernel, mbots_profile): python tools/shift_time.py WORLDS
MBOTS_SHIFT_BULK=1 selects the bulk-copy variant (csrc/shift_bulk.cu)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "madrona-bots_b200"))
import torch
import bench
from madrona_bots import SimManager
W = int(sys.argv[1])
sim = SimManager(0, W, 0, 32)
pool = [torch.from_numpy(a).cuda() for a in bench.one_hot_pool_numpy(W * 48, 1234)]
for i in range(10):
    sim.write_actions(pool[i % 4]); sim.step(); sim.shift_observations()
sim.profile(True)
for i in range(40):
    sim.write_actions(pool[i % 4]); sim.step_async(); sim.shift_observations_async(); sim.sync()
torch.cuda.synchronize()
p = sim.profile_read()
A = sim.num_agents()
ms = p["shift"]["ms"] / max(1, p["shift"]["calls"])
print(json.dumps({"worlds": W, "bulk": os.environ.get("MBOTS_SHIFT_BULK", "0"), "agents": A, "shift_ms": round(ms, 5),
                  "GBps": round(264 * A / (ms * 1e-3) / 1e9, 1), "calls": p["shift"]["calls"]}))
