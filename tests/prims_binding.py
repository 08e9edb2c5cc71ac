"""Builds tests/prims_shim.cpp (the product's arithmetic header compiled for the host) -- test infrastructure."""
import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.path.join(HERE, "prims_shim.cpp")
OUT = os.path.join(HERE, "_build", "libprims_shim.so")
DEPS = [SRC, os.path.join(ROOT, "madrona-bots_b200", "csrc", "sim_prims.cuh"),
        os.path.join(ROOT, "include", "mbots_detmath.h")]

_lib = None


def shim():
    global _lib
    if _lib is not None:
        return _lib
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in DEPS):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-fno-fast-math",
                               "-march=x86-64-v3", "-x", "c++", SRC, "-o", OUT])
    L = ctypes.CDLL(OUT)
    u32, i32, f32 = ctypes.c_uint32, ctypes.c_int32, ctypes.c_float
    pf, pu = ctypes.POINTER(f32), ctypes.POINTER(u32)
    L.shim_split.argtypes = [u32, u32, u32, u32, pu]
    L.shim_init_key.argtypes = [u32, pu]
    L.shim_sample_i32.argtypes = [u32, u32, i32, i32]; L.shim_sample_i32.restype = i32
    L.shim_sample_uniform.argtypes = [u32, u32]; L.shim_sample_uniform.restype = f32
    L.shim_turn.argtypes = [f32, f32, ctypes.c_int, pf]
    L.shim_forward.argtypes = [f32, f32, pf]
    L.shim_normalize.argtypes = [f32, f32, pf]
    L.shim_move.argtypes = [pf, pf, pf, pf] + [ctypes.c_int] * 4; L.shim_move.restype = f32
    L.shim_chunk_of.argtypes = [f32, f32]; L.shim_chunk_of.restype = ctypes.c_int
    L.shim_cell_of.argtypes = [f32, f32]; L.shim_cell_of.restype = u32
    L.shim_surroundings.argtypes = [f32, f32, pu, pu, pf]
    L.shim_species_reward.argtypes = [u32, u32, u32]; L.shim_species_reward.restype = f32
    L.shim_agent_reward.argtypes = [f32, i32, ctypes.c_int, ctypes.c_int, ctypes.c_int]; L.shim_agent_reward.restype = f32
    L.shim_view_yaw.argtypes = [f32, f32]; L.shim_view_yaw.restype = f32
    L.shim_cull_trials.argtypes = [ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                   ctypes.POINTER(ctypes.c_longlong)]
    L.shim_cull_trials.restype = ctypes.c_longlong
    L.shim_wedge_trials.argtypes = [ctypes.c_uint64, ctypes.c_longlong, ctypes.POINTER(ctypes.c_longlong)]
    L.shim_wedge_trials.restype = ctypes.c_longlong
    _lib = L
    return L
