"""Drop-in for the reference's pybind module ``lunar_lander_cpp`` (src/lunar_lander_cpp.cpp:95-110): the classes
``GameLogicCppDouble`` / ``GameLogicCppFloat`` and ``generate_random_pad_positions`` with the same names, arguments,
attributes and exceptions, so that ``main.py`` / a headless driver can ``import lunar_lander_cpp`` unchanged.

The lander itself lives on the GPU: every method is one call of the ``ll_sim_*`` entry points of include/ll_gpu.h, which run
the device functions of the episode kernel on a single lander (csrc/ll_sim.cu).  Nothing is simulated on the CPU; without the
built library or a CUDA device construction fails.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Tuple

import numpy as np

from . import _abi
from ._abi import check, lib
from .config import Config

_ATTRS = ("x", "y", "vx", "vy", "fuel", "landing_pad_center_x")
_FLAGS = ("landed", "crashed", "landed_successfully")


class _GameLogic:
    """GameLogicCpp<T> (src/game_logic.h:15-93).  ``config_file`` empty = the process-wide configuration already loaded
    (module attribute ``default_config``), like the reference's constructor (src/game_logic.cpp:33-46)."""
    _precision = 64
    _dtype = np.float64

    def __init__(self, no_print_flag: bool = False, config_file: str = ""):
        global default_config
        if config_file:
            cfg = Config()
            if not cfg.loadConfiguration(config_file):
                # src/game_logic.cpp:42-45
                raise RuntimeError(f"GameLogicCpp: Failed to load configuration from {config_file}")
            default_config = cfg
        elif default_config is None:
            raise RuntimeError("GameLogicCpp: no configuration loaded (pass config_file)")
        self._cfg = default_config
        self.no_print = bool(no_print_flag)
        sizes = (C.c_uint64 * 2)(5, 4)  # the context needs a net shape; the stepped lander does not use it
        net = _abi.NetCfg()
        net.layer_sizes = C.cast(sizes, C.POINTER(C.c_uint64))
        net.n_layers, net.activation_id, net.population, net.top, net.seed = 2, _abi.LL_ACT_TANH, 1, 1, 0
        net.precision, net.mode, net.rank, net.world, net.device = self._precision, _abi.LL_MODE_EXACT, 0, 1, -1
        env = self._cfg.env_cfg()
        self._h = C.c_void_p()
        check(lib.ll_create(C.byref(self._h), C.byref(env), C.byref(net)))
        self._v = _abi.SimView()
        check(lib.ll_sim_reset(self._h, 0, 0.0, 0.0, C.byref(self._v)), self._h)

    def __del__(self):
        try:
            if self._h:
                lib.ll_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    # ---- def_readwrite members (src/lunar_lander_cpp.cpp:27-38)
    def _push(self, name, value):
        v = _abi.SimView()
        C.memmove(C.byref(v), C.byref(self._v), C.sizeof(v))
        setattr(v, name, value)
        check(lib.ll_sim_set(self._h, C.byref(v), C.byref(self._v)), self._h)

    landing_pad_y = property(lambda self: self._v.landing_pad_y)
    time_step = property(lambda self: int(self._v.time_step))

    # ---- methods
    def reset(self, spad_x1: Optional[float] = None, lpad_x1_new: Optional[float] = None) -> None:
        if (spad_x1 is None) != (lpad_x1_new is None):
            raise TypeError("reset(): incompatible function arguments (expected () or (spad_x1, lpad_x1_new))")
        if spad_x1 is None:
            check(lib.ll_sim_reset(self._h, 0, 0.0, 0.0, C.byref(self._v)), self._h)
        else:
            check(lib.ll_sim_reset(self._h, 1, float(spad_x1), float(lpad_x1_new), C.byref(self._v)), self._h)

    def is_done(self) -> bool:
        return bool(self._v.done)

    def _out(self, arr, who):
        if not isinstance(arr, np.ndarray) or arr.dtype != self._dtype:
            raise TypeError(f"{who}: incompatible function arguments (expected a numpy array of {np.dtype(self._dtype).name})")
        if arr.ndim != 1:
            raise RuntimeError(f"{who}: Output {'state ' if who == 'update' else ''}array must be 1-dimensional.")
        if arr.shape[0] < 5:
            raise AssertionError("state buffer must hold at least 5 values")  # assert, src/game_logic.cpp:277
        arr[:5] = np.frombuffer(self._v.state, dtype=np.float64, count=5).astype(self._dtype)

    def get_state(self, state_output: np.ndarray) -> None:
        self._out(state_output, "get_state")

    def update(self, action: int, state_output: np.ndarray) -> bool:
        was_done = bool(self._v.done)
        check(lib.ll_sim_update(self._h, int(action), C.byref(self._v)), self._h)
        self._out(state_output, "update")
        if not self.no_print and not was_done and self._v.done:  # src/game_logic.cpp:216, 224-238
            if self._v.landed_successfully:
                print("GameLogicCpp: Successful landing!")
            else:
                # the speed messages of the reference can never fire: vx, vy are zeroed before they are tested (:221-237)
                e = self._cfg.env_cfg()
                lpad_x1 = self._v.landing_pad_center_x - e.lpad_width / 2.0
                if not (self._v.x >= lpad_x1 and self._v.x + e.lander_width <= lpad_x1 + e.lpad_width):
                    print("GameLogicCpp: Crashed - outside landing area!")
        return bool(self._v.done)

    def calculate_step_penalty(self, action: int) -> float:
        p = C.c_double()
        check(lib.ll_sim_step_penalty(self._h, int(action), C.byref(p)), self._h)
        return p.value

    def calculate_terminal_penalty(self, steps_taken: int) -> float:
        p = C.c_double()
        check(lib.ll_sim_terminal_penalty(self._h, int(steps_taken), -1, C.byref(p), None), self._h)
        return p.value

    def calculate_terminal_penalty_lr_bonus(self, steps_taken: int, direction: int) -> Tuple[float, list]:
        if direction not in (0, 1):
            raise AssertionError("Direction must be 0 or 1")  # assert, src/game_logic.cpp:356
        p, bonus = C.c_double(), (C.c_double * 2)(0.0, 0.0)
        check(lib.ll_sim_terminal_penalty(self._h, int(steps_taken), int(direction), C.byref(p), bonus), self._h)
        return p.value, [bonus[0], bonus[1]]


def _attr(name, cast):
    return property(lambda self: cast(getattr(self._v, name)), lambda self, value: self._push(name, cast(value)))


for _n in _ATTRS:
    setattr(_GameLogic, _n, _attr(_n, float))
for _n in _FLAGS:
    setattr(_GameLogic, _n, _attr(_n, bool))
_GameLogic.last_action = _attr("last_action", int)


class GameLogicCppDouble(_GameLogic):
    _precision, _dtype = 64, np.float64


class GameLogicCppFloat(_GameLogic):
    _precision, _dtype = 32, np.float32


default_config: Optional[Config] = None
if os.path.exists("config.txt"):  # the reference's Python side loads ./config.txt at import (mod_config.py)
    _c = Config()
    if _c.loadConfiguration("config.txt"):
        default_config = _c


def generate_random_pad_positions(screen_width: float, spad_width: int, lpad_width: int, seed: Optional[int] = None,
                                  force_left_to_right: bool = False, force_right_to_left: bool = False) -> Tuple[int, int]:
    """src/game_utils.cpp:12-74 (std::mt19937 + uniform_int_distribution, rejection of overlapping pads, optional mirroring)."""
    s, l = C.c_int32(), C.c_int32()
    check(lib.ll_generate_random_pad_positions(float(screen_width), int(spad_width), int(lpad_width), 0 if seed is None else 1,
                                               0 if seed is None else int(seed), int(bool(force_left_to_right)),
                                               int(bool(force_right_to_left)), C.byref(s), C.byref(l)))
    return s.value, l.value
