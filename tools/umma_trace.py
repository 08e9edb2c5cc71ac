#!/usr/bin/env python
"""clock64 timeline of CTA 0 of policy_umma_kernel (library built with -DMBK_UMMA_TRACE):
   MBOTS_LIB=.../libmbots_trace.so python tools/umma_trace.py [rtype]"""
import ctypes, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "madrona-bots_b200"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests"))
import numpy as np
import torch
from madrona_bots import SimManager, _native
from madrona_bots.policy import FusedPolicy, SpeciesPolicy
from test_policy import _make

rtype = sys.argv[1] if len(sys.argv) > 1 else "GRU"
net, config = _make(rtype, ["Tanh", "ReLU"])
fused = FusedPolicy(SpeciesPolicy(config, net.state_dict(), "cuda:0"))
sim = SimManager(0, 16384, 0, 32)
sim.step()
A = sim.num_agents()
lib = _native.lib()
buf = np.zeros(5 * 2 * 2048, dtype=np.uint64)
for rep in range(3):
    lib.mbots_debug_umma_trace(None, 1)
    sim.policy_act(0, A, fused, 0, rep)
lib.mbots_debug_umma_trace(buf.ctypes.data_as(ctypes.c_void_p), 0)
ev = buf.reshape(5 * 2048, 2)
actor = np.repeat(np.arange(5), 2048)
keep = (ev[:, 0] >> 31) & 1 == 1
ev, actor = ev[keep], actor[keep]
order = np.argsort(ev[:, 1], kind="stable")
t0 = int(ev[order[0], 1])
names = {1: "epi wake", 3: "epi math done", 5: "epi fenced", 10: "iss w_full", 11: "iss a_ready", 12: "iss issued", 13: "iss w_empty"}
who = ["issuer", "epi t0 h0", "epi t0 h1", "epi t1 h0", "epi t1 h1"]
print("agents", A, "events", len(ev))
first = int(sys.argv[3]) if len(sys.argv) > 3 else 0          # skip the ramp: a window of the steady state
for i in order[first:first + (int(sys.argv[2]) if len(sys.argv) > 2 else 300)]:
    code = int(ev[i, 0]) & 0x7fffffff
    print(f"{int(ev[i, 1]) - t0:9d}  {who[actor[i]]:10s} tile {(code >> 8) & 255} block {code >> 16:2d}  {names.get(code & 255, code & 255)}")
