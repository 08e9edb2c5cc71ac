"""ctypes binding of oracle/_ref/librefsim.so -- ONE world of the reference, run from the reference's own source
text (src/sim/sim.cpp, M/src/render/ecs_system.cpp, M/src/mw/device/bvh_raycast.cpp) under the host mock of
oracle/refmock/. Test infrastructure: used by tests/test_reference_bodies.py to pin the CPU oracle."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
FIX_U5, FIX_U6, FIX_U19, FIX_U8, FIX_D3 = 1, 2, 4, 8, 16
CANONICAL = FIX_U5 | FIX_U6 | FIX_U19 | FIX_U8 | FIX_D3

_LIBS = {}


def refsim_lib(libm=False):
    """The library built with the shared transcendental definitions (rule D2) or, libm=True, with the host's
    cosf / sinf / atan2f. (Re)built where /root/reference is mounted; otherwise the shipped .so is used."""
    name = "librefsim_libm.so" if libm else "librefsim.so"
    if name in _LIBS:
        return _LIBS[name]
    path = os.path.join(REF_DIR, name)
    if os.path.isdir("/root/reference/src/sim"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "refsim"], stdout=subprocess.DEVNULL)
    if not os.path.exists(path):
        return None
    lib = ctypes.CDLL(path)
    lib.mbr_create.restype = ctypes.c_void_p
    lib.mbr_create.argtypes = [ctypes.c_uint32] * 4
    for f in ("mbr_destroy", "mbr_step", "mbr_shift"):
        getattr(lib, f).argtypes = [ctypes.c_void_p]
        getattr(lib, f).restype = None
    lib.mbr_num_agents.argtypes = [ctypes.c_void_p]
    lib.mbr_num_obs_rows.argtypes = [ctypes.c_void_p]
    lib.mbr_set_actions.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32]
    lib.mbr_set_hidden.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32]
    lib.mbr_agents.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.mbr_obs.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.mbr_world.argtypes = [ctypes.c_void_p] * 5
    lib.mbr_ray_sphere.argtypes = [ctypes.c_void_p, ctypes.c_float, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    _LIBS[name] = lib
    return lib


class RefWorld:
    def __init__(self, seed, world_id, init_agents, flags=CANONICAL, libm=False):
        self.lib = refsim_lib(libm)
        if self.lib is None:
            raise RuntimeError("oracle/_ref/librefsim.so unavailable")
        self.h = ctypes.c_void_p(self.lib.mbr_create(seed, world_id, init_agents, flags))

    def __del__(self):
        if getattr(self, "h", None) and self.h.value:
            self.lib.mbr_destroy(self.h)
            self.h = ctypes.c_void_p()

    def step(self):
        self.lib.mbr_step(self.h)

    def shift_observations(self):
        self.lib.mbr_shift(self.h)

    def num_agents(self):
        return self.lib.mbr_num_agents(self.h)

    def num_obs_rows(self):
        return self.lib.mbr_num_obs_rows(self.h)

    def set_actions(self, a):
        a = np.ascontiguousarray(a, dtype=np.int32)
        self.lib.mbr_set_actions(self.h, a.ctypes.data, a.shape[0])

    def set_hidden(self, v):
        v = np.ascontiguousarray(v, dtype=np.float32)
        self.lib.mbr_set_hidden(self.h, v.ctypes.data, v.shape[0])

    def agents(self):
        n = self.num_agents()
        pose = np.zeros((max(n, 1), 8), dtype=np.float32)
        ints = np.zeros((max(n, 1), 4), dtype=np.int32)
        self.lib.mbr_agents(self.h, pose.ctypes.data, ints.ctypes.data)
        return pose[:n], ints[:n]

    def obs(self, prev=False):
        n = self.num_obs_rows()
        f = np.zeros((max(n, 1), 32), dtype=np.float32)
        i = np.zeros((max(n, 1), 10), dtype=np.int32)
        b = np.zeros((max(n, 1), 64), dtype=np.uint8)
        self.lib.mbr_obs(self.h, 1 if prev else 0, f.ctypes.data, i.ctypes.data, b.ctypes.data)
        return f[:n], i[:n], b[:n]

    def world(self):
        oi = np.zeros(8, dtype=np.int32)
        of = np.zeros(4, dtype=np.float32)
        food = np.zeros((48, 8), dtype=np.uint8)
        chunks = np.zeros((48, 2), dtype=np.uint32)
        self.lib.mbr_world(self.h, oi.ctypes.data, of.ctypes.data, food.ctypes.data, chunks.ctypes.data)
        return oi, of, food, chunks


# exports of the oracle / CUDA path that one reference world can be asked for (the semantic channel has no source
# at the pinned commit -- SURVEY.md section 8a -- and DEBUG_ROW_BASE / episode statistics are artefacts of this build)
REF_EXPORTS = [
    "ACTION", "PREV_ACTION", "HIDDEN_STATE", "PREV_HIDDEN_STATE", "REWARD", "PREV_REWARD", "POSITION",
    "PREV_POSITION", "HEALTH", "PREV_HEALTH", "SURROUNDING", "PREV_SURROUNDING", "SENSOR_DEPTH",
    "PREV_SENSOR_DEPTH", "STATS", "PREV_STATS", "SENSOR_INDEX", "SPECIES_COUNT", "DEBUG_WORLD_COUNT",
    "DEBUG_RNG_COUNT", "DEBUG_NUM_FOOD", "DEBUG_FOOD", "DEBUG_POSE", "DEBUG_AGENT_HEALTH", "DEBUG_AGENT_SPECIES",
    "DEBUG_AGENT_TARGET",
]


class RefSimAdapter:
    """One reference world behind the host-array surface of tests/oracle_binding.OracleSim (`tensor(name)`,
    `write(name, rows)`, `step`, `shift_observations`, `num_agents`), in the oracle's export formats, so the same
    driver and the same digests serve the reference text, the oracle and the CUDA path."""

    def __init__(self, seed, world_id, init_agents, flags=CANONICAL, libm=False):
        self.w = RefWorld(seed, world_id, init_agents, flags, libm)
        self.stepped = False

    def num_agents(self):
        return self.w.num_agents()

    def _row_of_export_row(self):
        """Before the first step the reference's observation rows are in creation order; the oracle's are already
        species-major. Export row r of the oracle belongs to the agent perm[r]."""
        pose, ints = self.w.agents()
        perm = np.argsort(ints[:, 1], kind="stable")
        return ints[perm, 3]

    def write(self, name, value):
        if not len(value):
            return
        rows = np.arange(len(value)) if self.stepped else self._row_of_export_row()
        buf = np.zeros_like(value)
        buf[rows] = value
        if name == "ACTION":
            self.w.set_actions(buf)
        elif name == "HIDDEN_STATE":
            self.w.set_hidden(buf)
        else:
            raise KeyError(name)

    def step(self):
        self.w.step()
        self.stepped = True

    def shift_observations(self):
        self.w.shift_observations()

    def tensor(self, name):
        assert self.stepped or name.startswith("DEBUG_") or name == "SPECIES_COUNT", "row exports are valid after the first step"
        w = self.w
        if name.startswith("DEBUG_AGENT") or name == "DEBUG_POSE" or name == "SENSOR_INDEX":
            pose, ints = w.agents()
            if name == "DEBUG_POSE":
                return np.ascontiguousarray(pose[:, [0, 1, 3, 6]])
            col = {"DEBUG_AGENT_HEALTH": 0, "DEBUG_AGENT_SPECIES": 1, "DEBUG_AGENT_TARGET": 2, "SENSOR_INDEX": 3}[name]
            return np.ascontiguousarray(ints[:, col:col + 1])
        if name in ("SPECIES_COUNT", "DEBUG_WORLD_COUNT", "DEBUG_RNG_COUNT", "DEBUG_NUM_FOOD", "DEBUG_FOOD"):
            oi, of, food, chunks = w.world()
            return {"SPECIES_COUNT": oi[None, :4].copy(), "DEBUG_WORLD_COUNT": oi[None, 6:7].copy(),
                    "DEBUG_RNG_COUNT": oi[None, 5:6].copy(), "DEBUG_NUM_FOOD": oi[None, 4:5].copy(), "DEBUG_FOOD": food}[name]
        prev = name.startswith("PREV_")
        f, iv, b = w.obs(prev)
        base = name[5:] if prev else name
        return np.ascontiguousarray({
            "POSITION": f[:, 0:2], "HEALTH": f[:, 2:3], "SURROUNDING": f[:, 3:5], "REWARD": f[:, 5:6],
            "HIDDEN_STATE": f[:, 7:23], "ACTION": iv[:, 0:6], "STATS": iv[:, 6:10], "SENSOR_DEPTH": b[:, 0:32],
        }[base])
