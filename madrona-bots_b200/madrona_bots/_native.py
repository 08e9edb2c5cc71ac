"""ctypes binding of libmbots.so (include/mbots.h). No fallback: if the library or a CUDA
device is missing, loading / creating a simulator raises."""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# MBOTS_LIB selects another build of the same library (A/B runs of kernel variants); default = in-tree build
LIB_PATH = os.environ.get("MBOTS_LIB") or os.path.join(os.path.dirname(_HERE), "lib", "libmbots.so")

# mbots_export_id
EXPORT = dict(
    RESET=0, ACTION=1, PREV_ACTION=2, HIDDEN_STATE=3, PREV_HIDDEN_STATE=4, REWARD=5, PREV_REWARD=6,
    DONE=7, POSITION=8, PREV_POSITION=9, HEALTH=10, PREV_HEALTH=11, SURROUNDING=12,
    PREV_SURROUNDING=13, SENSOR_SEMANTIC=14, SENSOR_DEPTH=15, PREV_SENSOR_SEMANTIC=16,
    PREV_SENSOR_DEPTH=17, STATS=18, PREV_STATS=19, SENSOR_INDEX=20, SPECIES_COUNT=21,
    DEBUG_WORLD_COUNT=100, DEBUG_WORLD_OFFSET=101, DEBUG_RNG_COUNT=102, DEBUG_NUM_FOOD=103,
    DEBUG_FOOD=104, DEBUG_POSE=105, DEBUG_AGENT_HEALTH=106, DEBUG_AGENT_SPECIES=107,
    DEBUG_AGENT_TARGET=108, DEBUG_ROW_BASE=109, DEBUG_ROW_OWNER=110,
)

# every symbol include/mbots.h declares
SYMBOLS = [
    "mbots_create", "mbots_create_ex", "mbots_destroy", "mbots_step", "mbots_step_async", "mbots_sync",
    "mbots_shift_observations", "mbots_shift_observations_async", "mbots_num_agents", "mbots_tensor",
    "mbots_set_action", "mbots_agent_offset_for_world", "mbots_copy_from_host", "mbots_copy_to_host",
    "mbots_episode_stats", "mbots_launch_count", "mbots_profile", "mbots_profile_read",
    "mbots_row_capacity", "mbots_pack_observations", "mbots_sample_actions", "mbots_policy_act", "mbots_state_bytes", "mbots_save_state", "mbots_load_state",
    "mbots_rollout_step_async", "mbots_write_actions", "mbots_graph_launch_count", "mbots_debug_stale_views", "mbots_species_blocks",
    "mbots_row_scan_static", "mbots_policy_act_species", "mbots_policy_act_all", "mbots_last_error", "mbots_version",
]


class PolicyOp(ctypes.Structure):          # mbots_policy_op (include/mbots_policy.h)
    _fields_ = [("op", ctypes.c_int32), ("act", ctypes.c_int32), ("k_pad", ctypes.c_uint32),
                ("w_ofs", ctypes.c_uint32), ("b_ofs", ctypes.c_uint32)]


class PolicyProgram(ctypes.Structure):     # mbots_policy_program
    _fields_ = [("n_ops", ctypes.c_int32), ("layout", ctypes.c_int32), ("ops", PolicyOp * 12)]


OP_LINEAR, OP_LSTM0, OP_GRU0, OP_HEAD = 0, 1, 2, 3
LAYOUT_ROWPAD, LAYOUT_SW128 = 0, 1
ACT = dict(NONE=0, Tanh=1, ELU=2, LogSigmoid=3, LeakyReLU=4, ReLU=5, Sigmoid=6)


class Config(ctypes.Structure):
    _fields_ = [("gpu_id", ctypes.c_int32), ("num_worlds", ctypes.c_uint32), ("rand_seed", ctypes.c_uint32),
                ("init_num_agents_per_world", ctypes.c_uint32), ("max_agents_per_world", ctypes.c_uint32),
                ("world_id_offset", ctypes.c_uint32)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C madrona-bots_b200/csrc` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
    L = ctypes.CDLL(LIB_PATH)
    vp, i32, u32, u64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_uint32, ctypes.c_uint64
    L.mbots_create.argtypes = [i32, u32, u32, u32, ctypes.POINTER(vp)]
    L.mbots_create_ex.argtypes = [ctypes.POINTER(Config), ctypes.POINTER(vp)]
    L.mbots_destroy.argtypes = [vp]
    for name in ("mbots_step", "mbots_step_async", "mbots_sync", "mbots_shift_observations",
                 "mbots_shift_observations_async"):
        getattr(L, name).argtypes = [vp, vp]
    L.mbots_rollout_step_async.argtypes = [vp, vp, u64, i32, vp]
    L.mbots_write_actions.argtypes = [vp, vp, u64, vp]
    L.mbots_graph_launch_count.argtypes = [vp, ctypes.POINTER(u64)]
    L.mbots_debug_stale_views.argtypes = [vp, i32]
    L.mbots_species_blocks.argtypes = [vp, ctypes.POINTER(u32)]
    L.mbots_row_scan_static.argtypes = [vp, u32, vp, vp, vp, vp]
    L.mbots_num_agents.argtypes = [vp, ctypes.POINTER(u32)]
    L.mbots_tensor.argtypes = [vp, i32, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(i32)]
    L.mbots_set_action.argtypes = [vp, u32, i32, i32, i32, i32, i32, i32]
    L.mbots_agent_offset_for_world.argtypes = [vp, u32, ctypes.POINTER(u32)]
    L.mbots_copy_from_host.argtypes = [vp, i32, vp, u64, vp]
    L.mbots_copy_to_host.argtypes = [vp, i32, vp, u64, vp]
    L.mbots_episode_stats.argtypes = [vp, ctypes.POINTER(ctypes.c_int64), vp]
    L.mbots_launch_count.argtypes = [vp, ctypes.POINTER(u64)]
    L.mbots_profile.argtypes = [vp, i32]
    L.mbots_profile_read.argtypes = [vp, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(u32)]
    L.mbots_pack_observations.argtypes = [vp, i32, u32, u32, vp, vp]
    L.mbots_sample_actions.argtypes = [vp, u32, u32, vp, u32, u32, vp, vp]
    L.mbots_policy_act.argtypes = [vp, u32, u32, vp, u64, ctypes.POINTER(PolicyProgram), u32, u32, i32, vp, vp, vp]
    L.mbots_policy_act_species.argtypes = [vp, i32, vp, u64, ctypes.POINTER(PolicyProgram), u32, u32, i32, vp, vp, vp]
    L.mbots_policy_act_all.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(u64), ctypes.POINTER(ctypes.POINTER(PolicyProgram)), u32, u32,
                                       i32, vp, vp, vp]
    L.mbots_row_capacity.argtypes = [vp, ctypes.POINTER(u64)]
    L.mbots_state_bytes.argtypes = [vp, ctypes.POINTER(u64)]
    L.mbots_save_state.argtypes = [vp, vp, u64, vp]
    L.mbots_load_state.argtypes = [vp, vp, u64, vp]
    L.mbots_last_error.restype = ctypes.c_char_p
    L.mbots_version.restype = ctypes.c_char_p
    for name in SYMBOLS:
        if name not in ("mbots_last_error", "mbots_version"):
            getattr(L, name).restype = ctypes.c_int
    _lib = L
    return L


def check(rc, what):
    if rc != 0:
        msg = lib().mbots_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed ({rc}): {msg}")
