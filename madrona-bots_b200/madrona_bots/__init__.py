"""madrona_bots -- drop-in replacement for the reference's nanobind module of the same name
(src/entry/entry.cpp:16-44), backed by libmbots.so (hand-written sm_100a kernels).

    from madrona_bots import SimManager
    sim = SimManager(gpu_id, num_worlds, rand_seed, init_num_agents_per_world)
    sim.step()
    obs = sim.depth_tensor(False).to_torch()      # zero-copy CUDA view, (A, 32) uint8

Same constructor, method names, tensor shapes and dtypes as the reference, so
learn/training_loop.py and learn/env.py run against it unchanged. Rows are species-major;
A (the leading dimension) is the live-agent count sampled when `*_tensor()` is called, so
callers re-fetch the tensors after every step, exactly as with the reference
(src/entry/mgr.cpp:70-76).
"""
import ctypes

from . import _native
from ._native import EXPORT

__all__ = ["SimManager", "Tensor", "ScriptBotsViewer"]

_TYPESTR = {0: "<i4", 1: "<f4", 2: "|i1", 3: "|u1"}
_ITEMSIZE = {0: 4, 1: 4, 2: 1, 3: 1}


class Tensor:
    """madrona.Tensor look-alike (M/src/python/bindings.cpp:52-68): a typed, shaped device
    pointer with `to_torch()` returning a zero-copy view on the simulator's GPU."""

    def __init__(self, ptr, shape, dtype_code, gpu_id, owner, row_indexed=True):
        self._row_indexed = row_indexed   # leading dimension = live agents (buffer holds `capacity` rows)
        self._ptr = ptr
        self._shape = tuple(int(x) for x in shape)
        self._code = dtype_code
        self._gpu_id = gpu_id
        self._owner = owner          # keeps the simulator (and its memory) alive

    @property
    def __cuda_array_interface__(self):
        return {"shape": self._shape, "typestr": _TYPESTR[self._code], "data": (self._ptr, False),
                "version": 2, "strides": None}

    def shape(self):
        return self._shape

    def devicePtr(self):
        return self._ptr

    def isOnGPU(self):
        return True

    def gpuID(self):
        return self._gpu_id

    def to_torch(self):
        import torch
        dt = {0: torch.int32, 1: torch.float32, 2: torch.int8, 3: torch.uint8}[self._code]
        dev = torch.device("cuda", self._gpu_id)
        if self._shape[0] == 0:
            return torch.empty(self._shape, dtype=dt, device=dev)
        # The library keeps every export at a fixed address, allocated for `capacity` rows: wrap the
        # pointer once per (address, dtype, columns) and re-slice that view to the live row count --
        # re-wrapping through __cuda_array_interface__ costs tens of microseconds per call.
        cache = getattr(self._owner, "_views", None)
        cap = getattr(self._owner, "_row_capacity", 0)
        if cache is None or not self._row_indexed or self._shape[0] > cap:
            t = torch.as_tensor(self, device=dev)
        else:
            key = (self._ptr, self._code, self._shape[1])
            base = cache.get(key)
            if base is None:
                base = torch.as_tensor(Tensor(self._ptr, (cap, self._shape[1]), self._code, self._gpu_id,
                                              None), device=dev)
                cache[key] = base
            t = base[: self._shape[0]]
            # the view handed out keeps the simulator (and so the device memory behind the pointer) alive; the cached
            # base does not, or the handle would sit in a reference cycle and be destroyed only by the cycle collector
            t._mbots_owner = self._owner
        assert t.data_ptr() == self._ptr, "expected a zero-copy view"
        return t


class SimManager:
    """Mirror of `Manager` as bound by the reference (src/entry/entry.cpp:16-44,
    src/entry/mgr.hpp:28-62)."""

    def __init__(self, gpu_id, num_worlds, rand_seed, init_num_agents_per_world, *,
                 max_agents_per_world=0, world_id_offset=0):
        self._lib = _native.lib()
        self._h = ctypes.c_void_p()
        self._gpu_id = int(gpu_id)
        self.num_worlds = int(num_worlds)
        cfg = _native.Config(int(gpu_id), int(num_worlds), int(rand_seed) & 0xFFFFFFFF,
                             int(init_num_agents_per_world), int(max_agents_per_world), int(world_id_offset))
        _native.check(self._lib.mbots_create_ex(ctypes.byref(cfg), ctypes.byref(self._h)), "mbots_create_ex")
        cap = ctypes.c_uint64()
        _native.check(self._lib.mbots_row_capacity(self._h, ctypes.byref(cap)), "mbots_row_capacity")
        self._row_capacity = int(cap.value)
        self._views = {}

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            self._lib.mbots_destroy(h)
            self._h = ctypes.c_void_p()

    def _stream(self):
        import torch
        return ctypes.c_void_p(torch.cuda.current_stream(self._gpu_id).cuda_stream)

    # ---- Manager::step / shiftObservations ------------------------------------------------
    def step(self):
        _native.check(self._lib.mbots_step(self._h, self._stream()), "mbots_step")

    def step_async(self):
        _native.check(self._lib.mbots_step_async(self._h, self._stream()), "mbots_step_async")

    def rollout_step_async(self, actions=None, shift=True):
        """One launch: [write `actions` (a contiguous CUDA int32 (rows, 6) tensor in observation-row order; only the live
        rows are written, the count is read on the device)] -> step + sensor -> [shift_observations]. Nothing is waited
        for: the host may enqueue the next step at once (`sync()` waits for the last one). `actions` is read when the
        step runs on the device: keep the tensor alive and unchanged until then."""
        ptr, rows = None, 0
        if actions is not None:
            assert actions.is_cuda and actions.is_contiguous() and actions.dim() == 2 and actions.shape[1] == 6
            ptr, rows = ctypes.c_void_p(actions.data_ptr()), int(actions.shape[0])
        _native.check(self._lib.mbots_rollout_step_async(self._h, ptr, rows, 1 if shift else 0, self._stream()),
                      "mbots_rollout_step_async")

    def write_actions(self, actions):
        assert actions.is_cuda and actions.is_contiguous() and actions.dim() == 2 and actions.shape[1] == 6
        _native.check(self._lib.mbots_write_actions(self._h, ctypes.c_void_p(actions.data_ptr()), int(actions.shape[0]),
                                                    self._stream()), "mbots_write_actions")

    def debug_stale_views(self, enable=True):
        """Debug mode for the pointer-lifetime rule (include/mbots.h): the next step raises if the action or hidden-state
        export was written through a view fetched before the previous step."""
        _native.check(self._lib.mbots_debug_stale_views(self._h, 1 if enable else 0), "mbots_debug_stale_views")

    def graph_launch_count(self):
        out = ctypes.c_uint64()
        _native.check(self._lib.mbots_graph_launch_count(self._h, ctypes.byref(out)), "mbots_graph_launch_count")
        return out.value

    def sync(self):
        _native.check(self._lib.mbots_sync(self._h, self._stream()), "mbots_sync")

    def shift_observations(self):
        _native.check(self._lib.mbots_shift_observations(self._h, self._stream()), "mbots_shift_observations")

    def shift_observations_async(self):
        _native.check(self._lib.mbots_shift_observations_async(self._h, self._stream()),
                      "mbots_shift_observations_async")

    # ---- exports --------------------------------------------------------------------------
    def export(self, name):
        ptr = ctypes.c_void_p()
        dims = (ctypes.c_int64 * 2)()
        dt = ctypes.c_int32()
        _native.check(self._lib.mbots_tensor(self._h, EXPORT[name], ctypes.byref(ptr), dims, ctypes.byref(dt)),
                      f"mbots_tensor({name})")
        per_world = name in ("RESET", "SPECIES_COUNT", "DEBUG_FOOD", "DEBUG_ROW_BASE") or name.startswith("DEBUG_WORLD") \
            or name in ("DEBUG_RNG_COUNT", "DEBUG_NUM_FOOD")
        return Tensor(ptr.value or 0, (dims[0], dims[1]), dt.value, self._gpu_id, self, row_indexed=not per_world)

    def _pair(self, cur, prev, is_prev):
        return self.export(prev if is_prev else cur)

    def depth_tensor(self, is_prev=False):
        return self._pair("SENSOR_DEPTH", "PREV_SENSOR_DEPTH", is_prev)

    def semantic_tensor(self, is_prev=False):
        return self._pair("SENSOR_SEMANTIC", "PREV_SENSOR_SEMANTIC", is_prev)

    def reward_tensor(self, is_prev=False):
        return self._pair("REWARD", "PREV_REWARD", is_prev)

    def species_count_tensor(self):
        return self.export("SPECIES_COUNT")

    def position_tensor(self, is_prev=False):
        return self._pair("POSITION", "PREV_POSITION", is_prev)

    def health_tensor(self, is_prev=False):
        return self._pair("HEALTH", "PREV_HEALTH", is_prev)

    def surrounding_tensor(self, is_prev=False):
        return self._pair("SURROUNDING", "PREV_SURROUNDING", is_prev)

    def action_tensor(self, is_prev=False):
        return self._pair("ACTION", "PREV_ACTION", is_prev)

    def stats_tensor(self, is_prev=False):
        return self._pair("STATS", "PREV_STATS", is_prev)

    def hidden_state_tensor(self, is_prev=False):
        return self._pair("HIDDEN_STATE", "PREV_HIDDEN_STATE", is_prev)

    # C++-only members of the reference's Manager (src/entry/mgr.cpp:241-277)
    def sensor_index_tensor(self):
        return self.export("SENSOR_INDEX")

    def set_action(self, agent_idx, forward, backward, rotate_left, rotate_right, shoot, breed):
        _native.check(self._lib.mbots_set_action(self._h, agent_idx, forward, backward, rotate_left,
                                                 rotate_right, shoot, breed), "mbots_set_action")

    def agent_offset_for_world(self, world_idx):
        out = ctypes.c_uint32()
        _native.check(self._lib.mbots_agent_offset_for_world(self._h, world_idx, ctypes.byref(out)),
                      "mbots_agent_offset_for_world")
        return out.value

    # ---- extensions -----------------------------------------------------------------------
    def num_agents(self):
        out = ctypes.c_uint32()
        _native.check(self._lib.mbots_num_agents(self._h, ctypes.byref(out)), "mbots_num_agents")
        return out.value

    def episode_stats(self):
        out = (ctypes.c_int64 * 16)()
        _native.check(self._lib.mbots_episode_stats(self._h, out, self._stream()), "mbots_episode_stats")
        return list(out)

    def launch_count(self):
        out = ctypes.c_uint64()
        _native.check(self._lib.mbots_launch_count(self._h, ctypes.byref(out)), "mbots_launch_count")
        return out.value

    def profile(self, enable):
        _native.check(self._lib.mbots_profile(self._h, 1 if enable else 0), "mbots_profile")

    def profile_read(self):
        ms = (ctypes.c_float * 4)()
        calls = (ctypes.c_uint32 * 4)()
        _native.check(self._lib.mbots_profile_read(self._h, ms, calls), "mbots_profile_read")
        names = ("step", "scan", "export_sense", "shift")
        return {n: {"ms": ms[i], "calls": calls[i]} for i, n in enumerate(names)}

    def copy_from_host(self, name, host_ptr, nbytes):
        _native.check(self._lib.mbots_copy_from_host(self._h, EXPORT[name], ctypes.c_void_p(host_ptr), nbytes,
                                                     self._stream()), "mbots_copy_from_host")

    def copy_to_host(self, name, host_ptr, nbytes, stream=None):
        """Asynchronous on `stream` (a torch.cuda.Stream; default: the current stream). The exports of a step stay
        intact until the step after next (double-buffered row sets), so a copy stream that waits on an event
        recorded after step t may drain them while step t + 1 runs -- as long as step t + 2 waits for the copy."""
        st = self._stream() if stream is None else ctypes.c_void_p(stream.cuda_stream)
        _native.check(self._lib.mbots_copy_to_host(self._h, EXPORT[name], ctypes.c_void_p(host_ptr), nbytes, st),
                      "mbots_copy_to_host")

    # ---- policy-inference ends --------------------------------
    def pack_observations(self, start, end, prev=False, out=None):
        """learn/util.py construct_obs(sim_mgr, start, end, prev) as one kernel: (end-start, 69) float32."""
        import torch
        n = int(end) - int(start)
        if out is None:
            out = torch.empty((n, 69), dtype=torch.float32, device=torch.device("cuda", self._gpu_id))
        assert out.is_contiguous() and out.dtype == torch.float32 and out.shape == (n, 69)
        _native.check(self._lib.mbots_pack_observations(self._h, 1 if prev else 0, int(start), int(end),
                                                        ctypes.c_void_p(out.data_ptr()), self._stream()),
                      "mbots_pack_observations")
        return out

    def species_blocks(self):
        """[(start, end)] * 4: the row range of every species in the current exports (rows are species-major,
        src/sim/sim.cpp:1148-1149). Exact, unlike the cumulative sum of `species_count_tensor` the reference's loop
        uses (learn/training_loop.py:43-45): those counts are taken before the step's respawns (sim.cpp:804-836)."""
        out = (ctypes.c_uint32 * 5)()                    # from the totals every step brings to the host: no device read
        _native.check(self._lib.mbots_species_blocks(self._h, out), "mbots_species_blocks")
        return [(int(out[k]), int(out[k + 1])) for k in range(4)]

    def sample_actions(self, start, end, weights, seed, draw, want_index=False):
        """Categorical sample of rows [start, end) from (n, 6) float32 action weights + one-hot write into
        the action export (training_loop.py:57,77-83,136). Returns the picked indices when asked to."""
        import torch
        n = int(end) - int(start)
        assert weights.is_contiguous() and weights.dtype == torch.float32 and weights.shape == (n, 6)
        idx = torch.empty(n, dtype=torch.int32, device=weights.device) if want_index else None
        _native.check(self._lib.mbots_sample_actions(self._h, int(start), int(end), ctypes.c_void_p(weights.data_ptr()),
                                                     int(seed) & 0xFFFFFFFF, int(draw) & 0xFFFFFFFF,
                                                     ctypes.c_void_p(idx.data_ptr()) if want_index else None,
                                                     self._stream()), "mbots_sample_actions")
        return idx

    def policy_act(self, start, end, fused, seed, draw, sample=True, want_index=False, want_logits=False):
        """One kernel: observations of rows [start, end) -> `fused` (madrona_bots.policy.FusedPolicy) -> sampled
        one-hot actions in the action export. Returns the picked indices, the logits, both, or None."""
        import torch
        n = int(end) - int(start)
        dev = torch.device("cuda", self._gpu_id)
        idx = torch.empty(n, dtype=torch.int32, device=dev) if want_index else None
        lg = torch.empty((n, 6), dtype=torch.float32, device=dev) if want_logits else None
        _native.check(self._lib.mbots_policy_act(
            self._h, int(start), int(end), ctypes.c_void_p(fused.blob.data_ptr()), fused.blob.numel(),
            ctypes.byref(fused.program), int(seed) & 0xFFFFFFFF, int(draw) & 0xFFFFFFFF, 1 if sample else 0,
            ctypes.c_void_p(lg.data_ptr()) if want_logits else None,
            ctypes.c_void_p(idx.data_ptr()) if want_index else None, self._stream()), "mbots_policy_act")
        if want_index and want_logits:
            return idx, lg
        return idx if want_index else lg

    def policy_act_species(self, species, fused, seed, draw, sample=True, out_index=None, out_logits=None):
        """`policy_act` for the whole block of one species (0..3), its row range resolved on the device: no host wait for
        the last step, valid while asynchronous steps are pending. `out_index` (rows,) int32 / `out_logits` (rows, 6)
        float32: optional caller-owned CUDA tensors indexed by observation row (size them for the population bound)."""
        _native.check(self._lib.mbots_policy_act_species(
            self._h, int(species), ctypes.c_void_p(fused.blob.data_ptr()), fused.blob.numel(), ctypes.byref(fused.program),
            int(seed) & 0xFFFFFFFF, int(draw) & 0xFFFFFFFF, 1 if sample else 0,
            ctypes.c_void_p(out_logits.data_ptr()) if out_logits is not None else None,
            ctypes.c_void_p(out_index.data_ptr()) if out_index is not None else None, self._stream()),
            "mbots_policy_act_species")

    def policy_act_all(self, fused4, seed, draw, sample=True, out_index=None, out_logits=None):
        """All four species in ONE kernel launch (`fused4`: four FusedPolicy in the sw128 layout, species order), row
        ranges resolved on the device: no host wait. Outputs as in `policy_act_species`."""
        blobs = (ctypes.c_void_p * 4)(*[f.blob.data_ptr() for f in fused4])
        sizes = (ctypes.c_uint64 * 4)(*[f.blob.numel() for f in fused4])
        progs = (ctypes.POINTER(_native.PolicyProgram) * 4)(*[ctypes.pointer(f.program) for f in fused4])
        _native.check(self._lib.mbots_policy_act_all(
            self._h, blobs, sizes, progs, int(seed) & 0xFFFFFFFF, int(draw) & 0xFFFFFFFF, 1 if sample else 0,
            ctypes.c_void_p(out_logits.data_ptr()) if out_logits is not None else None,
            ctypes.c_void_p(out_index.data_ptr()) if out_index is not None else None, self._stream()),
            "mbots_policy_act_all")

    # ---- state checkpoint / restore ---------------------------
    def save_state(self, path=None):
        """Every persistent array of the simulator as one uint8 numpy blob (also written to `path`
        when given). `load_state` of that blob into a handle of the same shape resumes the rollout
        bit-exactly."""
        import numpy as np
        n = ctypes.c_uint64()
        _native.check(self._lib.mbots_state_bytes(self._h, ctypes.byref(n)), "mbots_state_bytes")
        blob = np.empty(n.value, dtype=np.uint8)
        _native.check(self._lib.mbots_save_state(self._h, ctypes.c_void_p(blob.ctypes.data), n.value,
                                                 self._stream()), "mbots_save_state")
        if path is not None:
            blob.tofile(path)
        return blob

    def load_state(self, blob_or_path):
        import numpy as np
        blob = np.fromfile(blob_or_path, dtype=np.uint8) if isinstance(blob_or_path, (str, bytes)) or \
            hasattr(blob_or_path, "__fspath__") else np.ascontiguousarray(blob_or_path, dtype=np.uint8)
        _native.check(self._lib.mbots_load_state(self._h, ctypes.c_void_p(blob.ctypes.data), blob.nbytes,
                                                 self._stream()), "mbots_load_state")


class ScriptBotsViewer:
    """The reference's Vulkan viewer (src/gfx, src/entry/entry.cpp:46-80). Unavailable on this
    box: there is no graphics stack here or on the GPU machines (out of scope)."""

    def __init__(self, *args, **kwargs):
        raise RuntimeError("ScriptBotsViewer needs a Vulkan graphics stack: unavailable on this box")
