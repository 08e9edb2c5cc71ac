"""ctypes binding of the CPU oracle (oracle/libmbots_oracle.so) -- test infrastructure.

The oracle exposes the same export ids as include/mbots.h with HOST pointers;
`OracleSim.tensor(id)` returns a numpy view of the oracle's own array (valid
until the next step), so tests can write actions / hidden state in place
exactly as Python callers of the reference do through `to_torch()` views
(learn/training_loop.py:136-137).
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "libmbots_oracle.so")
ORACLE_REF_SO = os.path.join(ORACLE_DIR, "_ref", "libmbots_oracle_ref.so")

# mbots_export_id (include/mbots.h), mirrored by name
EXPORT = dict(
    RESET=0, ACTION=1, PREV_ACTION=2, HIDDEN_STATE=3, PREV_HIDDEN_STATE=4, REWARD=5, PREV_REWARD=6,
    DONE=7, POSITION=8, PREV_POSITION=9, HEALTH=10, PREV_HEALTH=11, SURROUNDING=12,
    PREV_SURROUNDING=13, SENSOR_SEMANTIC=14, SENSOR_DEPTH=15, PREV_SENSOR_SEMANTIC=16,
    PREV_SENSOR_DEPTH=17, STATS=18, PREV_STATS=19, SENSOR_INDEX=20, SPECIES_COUNT=21,
    DEBUG_WORLD_COUNT=100, DEBUG_WORLD_OFFSET=101, DEBUG_RNG_COUNT=102, DEBUG_NUM_FOOD=103,
    DEBUG_FOOD=104, DEBUG_POSE=105, DEBUG_AGENT_HEALTH=106, DEBUG_AGENT_SPECIES=107,
    DEBUG_AGENT_TARGET=108, DEBUG_ROW_BASE=109,
)
ROW_EXPORTS = [k for k, v in EXPORT.items() if 1 <= v <= 19 and k != "DONE"]
DTYPES = {0: np.int32, 1: np.float32, 2: np.int8, 3: np.uint8}


def build_oracle(ref=False):
    """(Re)build the oracle with its Makefile if the .so is missing or stale."""
    target = ORACLE_REF_SO if ref else ORACLE_SO
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("mbots_oracle.cpp", "oracle_prims.h")] + [
        os.path.join(ROOT, "include", f) for f in ("mbots_detmath.h", "mbots.h")]
    stale = not os.path.exists(target) or any(os.path.getmtime(s) > os.path.getmtime(target) for s in srcs)
    if stale:
        if ref and not os.path.isdir("/root/reference/external/madrona/include"):
            return None
        subprocess.check_call(["make", "-C", ORACLE_DIR, "ref" if ref else "all"],
                              stdout=subprocess.DEVNULL)
    return target


def _load(path):
    lib = ctypes.CDLL(path)
    lib.mbo_flavour.restype = ctypes.c_char_p
    lib.mbo_create.argtypes = [ctypes.c_uint32] * 5 + [ctypes.POINTER(ctypes.c_void_p)]
    lib.mbo_destroy.argtypes = [ctypes.c_void_p]
    lib.mbo_step.argtypes = [ctypes.c_void_p]
    lib.mbo_shift_observations.argtypes = [ctypes.c_void_p]
    lib.mbo_num_agents.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint32)]
    lib.mbo_tensor.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.POINTER(ctypes.c_void_p),
                               ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int32)]
    lib.mbo_episode_stats.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int64)]
    lib.mbo_debug_planarity.argtypes = [ctypes.c_void_p]
    lib.mbo_debug_planarity.restype = ctypes.c_float
    lib.mbo_set_threads.argtypes = [ctypes.c_int]
    lib.mbo_prim_sample_i32.argtypes = [ctypes.c_uint32, ctypes.c_uint32, ctypes.c_int32, ctypes.c_int32]
    lib.mbo_prim_sample_i32.restype = ctypes.c_int32
    lib.mbo_prim_sample_uniform.argtypes = [ctypes.c_uint32, ctypes.c_uint32]
    lib.mbo_prim_sample_uniform.restype = ctypes.c_float
    lib.mbo_prim_angle_axis_z.argtypes = [ctypes.c_float, ctypes.POINTER(ctypes.c_float)]
    return lib


_LIBS = {}


def oracle_lib(ref=False):
    key = "ref" if ref else "native"
    if key not in _LIBS:
        path = build_oracle(ref)
        if path is None or not os.path.exists(path):
            return None
        _LIBS[key] = _load(path)
    return _LIBS[key]


class OracleSim:
    """Same surface as madrona_bots.SimManager, numpy instead of torch."""

    def __init__(self, num_worlds, rand_seed, init_num_agents_per_world, cap=0, world_id_offset=0, ref=False):
        self.lib = oracle_lib(ref)
        if self.lib is None:
            raise RuntimeError("oracle library unavailable")
        self.h = ctypes.c_void_p()
        rc = self.lib.mbo_create(num_worlds, rand_seed, init_num_agents_per_world, cap, world_id_offset,
                                 ctypes.byref(self.h))
        if rc != 0:
            raise RuntimeError(f"mbo_create failed: {rc}")
        self.num_worlds = num_worlds

    def __del__(self):
        if getattr(self, "h", None) and self.h.value:
            self.lib.mbo_destroy(self.h)
            self.h = ctypes.c_void_p()

    def step(self):
        self.lib.mbo_step(self.h)

    def shift_observations(self):
        self.lib.mbo_shift_observations(self.h)

    def num_agents(self):
        n = ctypes.c_uint32()
        self.lib.mbo_num_agents(self.h, ctypes.byref(n))
        return n.value

    def tensor(self, name):
        ptr = ctypes.c_void_p()
        dims = (ctypes.c_int64 * 2)()
        dt = ctypes.c_int32()
        rc = self.lib.mbo_tensor(self.h, EXPORT[name], ctypes.byref(ptr), dims, ctypes.byref(dt))
        if rc != 0:
            raise RuntimeError(f"mbo_tensor({name}) failed: {rc}")
        dtype = np.dtype(DTYPES[dt.value])
        n = dims[0] * dims[1]
        if n == 0:
            return np.zeros((dims[0], dims[1]), dtype=dtype)
        buf = (ctypes.c_char * (n * dtype.itemsize)).from_address(ptr.value)
        return np.frombuffer(buf, dtype=dtype).reshape(dims[0], dims[1])

    def episode_stats(self):
        out = (ctypes.c_int64 * 16)()
        self.lib.mbo_episode_stats(self.h, out)
        return np.array(out[:], dtype=np.int64)

    def planarity(self):
        return float(self.lib.mbo_debug_planarity(self.h))


def random_actions(rng, n, mode="onehot"):
    """Action rows as learn/training_loop.py:82-83,136 writes them (one-hot int32), or
    independent Bernoulli(0.5) bits per field to exercise the if / else-if precedence of
    src/sim/sim.cpp:456-475."""
    if mode == "onehot":
        a = np.zeros((n, 6), dtype=np.int32)
        a[np.arange(n), rng.integers(0, 6, size=n)] = 1
        return a
    return rng.integers(0, 2, size=(n, 6)).astype(np.int32)
