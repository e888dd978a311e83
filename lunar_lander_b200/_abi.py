"""ctypes binding of libll_gpu.so — exactly the entry points declared in include/ll_gpu.h.

This is the only place the Python host side touches native code.  There is no fallback: if the shared library has not
been built (``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C lunar_lander_b200/csrc``) importing
this module raises, and without a CUDA device ``ll_create`` returns LL_ENODEVICE.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libll_gpu.so")

LL_OK, LL_EINVAL, LL_ENODEVICE, LL_ECUDA, LL_EIO, LL_ENOTOWNED = range(6)
LL_MODE_EXACT, LL_MODE_FAST = 0, 1
LL_ACT_SIGMOID, LL_ACT_TANH = 0, 1
LL_STRATEGY_ALL_FROM_ELITES, LL_STRATEGY_MIXED_WITH_RANDOM_INJECTION = 1, 2

ENV_FIELDS = [
    "pad_y1", "terrain_y", "max_vx", "max_vy", "g", "mu_x", "mu_y", "lander_width", "lander_height", "max_fuel",
    "lpad_x1", "lpad_width", "display_width", "display_height", "spad_width", "x0", "y0", "vx0", "vy0", "ax0", "ay0",
    "sp_dist_factor", "sp_action_factor", "tp_steps_factor", "tp_dist_factor", "tp_landed_bonus", "tp_landed_lr_bonus",
    "tp_fuel_bonus_factor", "tp_crashed_penalty", "tp_crash_v_mag_factor", "tp_no_fuel_penalty",
]


class EnvCfg(C.Structure):
    _fields_ = [(n, C.c_double) for n in ENV_FIELDS]


class NetCfg(C.Structure):
    _fields_ = [
        ("layer_sizes", C.POINTER(C.c_uint64)), ("n_layers", C.c_int32), ("activation_id", C.c_int32),
        ("population", C.c_uint64), ("top", C.c_uint64), ("seed", C.c_uint64), ("precision", C.c_int32),
        ("mode", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32), ("device", C.c_int32), ("reserved", C.c_int32),
        ("sigma_init", C.c_double), ("sigma_mut", C.c_double),
    ]


class GenStats(C.Structure):
    _fields_ = [("best_fitness", C.c_double), ("best_index", C.c_uint64), ("sum_fitness", C.c_double),
                ("sum_steps", C.c_double), ("best_lr", C.c_double), ("members", C.c_uint64)]


class StepRec(C.Structure):
    _fields_ = [("action", C.c_double), ("state", C.c_double * 5), ("x", C.c_double), ("y", C.c_double),
                ("vx", C.c_double), ("vy", C.c_double), ("fuel", C.c_double)]


class EpisodeResult(C.Structure):
    _fields_ = [("steps", C.c_int32), ("flags", C.c_int32), ("step_penalty_sum", C.c_double),
                ("terminal_penalty", C.c_double), ("bonus_lr", C.c_double * 2)]


class Cand(C.Structure):
    _fields_ = [("fitness", C.c_double), ("index", C.c_uint64)]


class Info(C.Structure):
    _fields_ = [("params", C.c_uint64), ("params_padded", C.c_uint64), ("pop_global", C.c_uint64),
                ("pop_local", C.c_uint64), ("first_global", C.c_uint64), ("top", C.c_uint64), ("precision", C.c_int32),
                ("mode", C.c_int32), ("rank", C.c_int32), ("world", C.c_int32), ("device", C.c_int32),
                ("sm_count", C.c_int32), ("device_name", C.c_char * 64), ("kernel", C.c_char * 64)]


class KernelTimes(C.Structure):
    _fields_ = [("episode_ms", C.c_double), ("episode_launches", C.c_uint64), ("select_ms", C.c_double),
                ("select_launches", C.c_uint64), ("gather_ms", C.c_double), ("gather_launches", C.c_uint64),
                ("mutate_ms", C.c_double), ("mutate_launches", C.c_uint64), ("other_launches", C.c_uint64)]


class SimView(C.Structure):
    _fields_ = [("x", C.c_double), ("y", C.c_double), ("vx", C.c_double), ("vy", C.c_double), ("fuel", C.c_double),
                ("landing_pad_center_x", C.c_double), ("landing_pad_y", C.c_double), ("landed", C.c_int32),
                ("crashed", C.c_int32), ("landed_successfully", C.c_int32), ("time_step", C.c_int32),
                ("last_action", C.c_int32), ("done", C.c_int32), ("state", C.c_double * 5)]


_ctx = C.c_void_p
_i32p, _u64p, _f64p = C.POINTER(C.c_int32), C.POINTER(C.c_uint64), C.POINTER(C.c_double)

# name -> (restype, argtypes); every symbol of include/ll_gpu.h
SIGNATURES = {
    "ll_abi_version": (C.c_int, []),
    "ll_last_error": (C.c_char_p, [_ctx]),
    "ll_create": (C.c_int, [C.POINTER(_ctx), C.POINTER(EnvCfg), C.POINTER(NetCfg)]),
    "ll_destroy": (None, [_ctx]),
    "ll_get_info": (C.c_int, [_ctx, C.POINTER(Info)]),
    "ll_set_stream": (C.c_int, [_ctx, C.c_void_p]),
    "ll_synchronize": (C.c_int, [_ctx]),
    "ll_set_population_strategy": (C.c_int, [_ctx, C.c_int, C.c_double]),
    "ll_create_population": (C.c_int, [_ctx, C.c_int]),
    "ll_set_sigmas": (C.c_int, [_ctx, C.c_double, C.c_double]),
    "ll_update_epochs": (C.c_int, [_ctx, C.c_uint64]),
    "ll_get_epochs": (C.c_int, [_ctx, _u64p]),
    "ll_set_epochs": (C.c_int, [_ctx, C.c_uint64]),
    "ll_select": (C.c_int, [_ctx, _u64p, C.c_uint64, _u64p]),
    "ll_local_topk": (C.c_int, [_ctx, C.POINTER(Cand), _u64p]),
    "ll_gather_elites": (C.c_int, [_ctx, _u64p, C.c_uint64, C.c_void_p]),
    "ll_set_elites": (C.c_int, [_ctx, C.c_void_p, C.c_uint64]),
    "ll_elite_buffer_bytes": (C.c_uint64, [_ctx]),
    "ll_comm_get_unique_id": (C.c_int, [C.c_void_p]),
    "ll_comm_init": (C.c_int, [_ctx, C.c_void_p]),
    "ll_comm_destroy": (C.c_int, [_ctx]),
    "ll_evaluate": (C.c_int, [_ctx, _i32p, _i32p, _u64p, C.c_int32, C.c_int32, _f64p, _f64p, _f64p, C.POINTER(GenStats)]),
    "ll_set_layouts": (C.c_int, [_ctx, _i32p, _i32p, _u64p, C.c_int32]),
    "ll_evaluate_device": (C.c_int, [_ctx, C.c_int32]),
    "ll_read_results": (C.c_int, [_ctx, _f64p, _f64p, _f64p, C.POINTER(GenStats)]),
    "ll_read_episodes": (C.c_int, [_ctx, _i32p, _f64p]),
    "ll_step_generation": (C.c_int, [_ctx, C.c_int32, C.c_int, C.POINTER(GenStats)]),
    "ll_trace_episode": (C.c_int, [_ctx, C.c_uint64, C.c_int32, C.c_int32, C.c_uint64, C.c_int32, C.POINTER(StepRec),
                                   C.c_int32, C.POINTER(EpisodeResult)]),
    "ll_feedforward_index": (C.c_int, [_ctx, C.c_void_p, C.c_uint64, _u64p]),
    "ll_get_member": (C.c_int, [_ctx, C.c_uint64, C.c_void_p]),
    "ll_set_member": (C.c_int, [_ctx, C.c_uint64, C.c_void_p]),
    "ll_set_members": (C.c_int, [_ctx, C.c_uint64, C.c_uint64, C.c_void_p]),
    "ll_get_members": (C.c_int, [_ctx, C.c_uint64, C.c_uint64, C.c_void_p]),
    "ll_mark_elites": (C.c_int, [_ctx]),
    "ll_serialize": (C.c_int, [_ctx, C.c_char_p, C.c_char_p, C.c_int32]),
    "ll_peek_weight_file": (C.c_int, [C.c_char_p, _u64p, C.c_int32, _i32p, _u64p, _u64p, _u64p, _i32p, _i32p, _f64p,
                                      C.c_char_p, C.c_int32]),
    "ll_deserialize": (C.c_int, [_ctx, C.c_char_p]),
    "ll_generate_random_pad_positions": (C.c_int, [C.c_double, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                                   C.c_int32, _i32p, _i32p]),
    "ll_balance_layouts": (C.c_int, [_i32p, _i32p, _f64p, _f64p, _u64p, C.c_int32, C.c_double, C.c_double, C.c_double,
                                     C.c_double]),
    "ll_sim_reset": (C.c_int, [_ctx, C.c_int32, C.c_double, C.c_double, C.POINTER(SimView)]),
    "ll_sim_update": (C.c_int, [_ctx, C.c_int32, C.POINTER(SimView)]),
    "ll_sim_get": (C.c_int, [_ctx, C.POINTER(SimView)]),
    "ll_sim_set": (C.c_int, [_ctx, C.POINTER(SimView), C.POINTER(SimView)]),
    "ll_sim_step_penalty": (C.c_int, [_ctx, C.c_int32, _f64p]),
    "ll_sim_terminal_penalty": (C.c_int, [_ctx, C.c_int32, C.c_int32, _f64p, _f64p]),
    "ll_profile": (C.c_int, [_ctx, C.c_int]),
    "ll_kernel_times_get": (C.c_int, [_ctx, C.POINTER(KernelTimes), C.c_int]),
    "ll_step_counter": (C.c_int, [_ctx, _f64p, _u64p, C.c_int]),
    "ll_microbench_fma": (C.c_int, [C.c_int32, C.c_int32, C.c_double, _f64p, _f64p]),
    "ll_microbench_copy": (C.c_int, [C.c_int32, C.c_uint64, C.c_int32, _f64p]),
    "ll_topk_static": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p]),
    "ll_tc05_probe": (C.c_int, [C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_float), _f64p]),
}


class LLError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libll_gpu error {code}: {msg}")
        self.code = code


def load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build the CUDA extension first (__graft_entry__.build() or "
            "`make -C lunar_lander_b200/csrc`). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype, fn.argtypes = res, args
    if lib.ll_abi_version() != 1:
        raise ImportError("libll_gpu.so ABI version mismatch")
    return lib


lib = load()


def check(rc: int, ctx=None) -> None:
    if rc != LL_OK:
        msg = lib.ll_last_error(ctx)
        raise LLError(rc, msg.decode() if msg else "?")
