"""Thin numpy-facing wrapper around one ``ll_ctx`` (one GPU shard of the population).

Everything here forwards 1:1 to include/ll_gpu.h; host buffers are numpy arrays, nothing is computed in Python.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import numpy as np

from . import _abi
from ._abi import check, lib
from .config import Config


def _p(a: np.ndarray, ty):
    return a.ctypes.data_as(C.POINTER(ty))


class Engine:
    def __init__(self, cfg: Config, sizes: Optional[Sequence[int]] = None, population: Optional[int] = None,
                 top: Optional[int] = None, precision: int = 64, mode: int = _abi.LL_MODE_EXACT, rank: int = 0,
                 world: int = 1, device: int = -1, seed: Optional[int] = None, sigma_init: float = 0.0,
                 sigma_mut: float = 0.0, activation: Optional[int] = None):
        self.cfg = cfg
        self.sizes = list(sizes) if sizes is not None else cfg.layer_sizes()
        self.precision = precision
        self.dtype = np.float64 if precision == 64 else np.float32
        self._sizes_arr = (C.c_uint64 * len(self.sizes))(*self.sizes)
        net = _abi.NetCfg()
        net.layer_sizes = C.cast(self._sizes_arr, C.POINTER(C.c_uint64))
        net.n_layers = len(self.sizes)
        net.activation_id = cfg.getInt("NNConfig.activation_id") if activation is None else activation
        net.population = cfg.getInt("NNConfig.population_size") if population is None else population
        net.top = cfg.getInt("NNConfig.top_individuals") if top is None else top
        net.seed = cfg.getInt("NNConfig.seed") if seed is None else seed
        net.precision, net.mode, net.rank, net.world, net.device = precision, mode, rank, world, device
        net.sigma_init, net.sigma_mut = sigma_init, sigma_mut
        env = cfg.env_cfg()
        self._h = C.c_void_p()
        check(lib.ll_create(C.byref(self._h), C.byref(env), C.byref(net)))
        info = _abi.Info()
        check(lib.ll_get_info(self._h, C.byref(info)), self._h)
        self.info = info
        self.P = int(info.params)
        self.pop_local, self.first_global, self.pop_global, self.top = (int(info.pop_local), int(info.first_global),
                                                                         int(info.pop_global), int(info.top))

    # ---- life cycle
    def close(self):
        if self._h:
            lib.ll_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream: int):
        check(lib.ll_set_stream(self._h, C.c_void_p(cuda_stream)), self._h)

    def synchronize(self):
        check(lib.ll_synchronize(self._h), self._h)

    # ---- GA
    def set_population_strategy(self, strategy: int, ratio: float):
        check(lib.ll_set_population_strategy(self._h, strategy, ratio), self._h)

    def set_sigmas(self, sigma_init: float, sigma_mut: float):
        check(lib.ll_set_sigmas(self._h, sigma_init, sigma_mut), self._h)

    def create_population(self, elitism: bool = True):
        check(lib.ll_create_population(self._h, int(elitism)), self._h)

    def update_epochs(self, n: int = 1):
        check(lib.ll_update_epochs(self._h, n), self._h)

    def get_epochs(self) -> int:
        v = C.c_uint64()
        check(lib.ll_get_epochs(self._h, C.byref(v)), self._h)
        return int(v.value)

    def set_epochs(self, n: int):
        check(lib.ll_set_epochs(self._h, n), self._h)

    def select(self, sorted_idx: Optional[np.ndarray] = None) -> np.ndarray:
        out = np.zeros(self.top, dtype=np.uint64)
        if sorted_idx is None:
            check(lib.ll_select(self._h, None, 0, _p(out, C.c_uint64)), self._h)
        else:
            s = np.ascontiguousarray(sorted_idx, dtype=np.uint64)
            check(lib.ll_select(self._h, _p(s, C.c_uint64), len(s), _p(out, C.c_uint64)), self._h)
        return out

    # ---- the exchange between GPUs, owned by the context (include/ll_gpu.h: ll_comm_*)
    @staticmethod
    def comm_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        check(lib.ll_comm_get_unique_id(buf))
        return buf.raw

    def comm_init(self, unique_id: bytes):
        if len(unique_id) != 128:
            raise ValueError("comm_init: the id is LL_COMM_ID_BYTES = 128 bytes")
        check(lib.ll_comm_init(self._h, C.c_char_p(unique_id)), self._h)

    def comm_destroy(self):
        check(lib.ll_comm_destroy(self._h), self._h)

    def local_topk(self) -> Tuple[np.ndarray, np.ndarray]:
        buf = (_abi.Cand * self.top)()
        n = C.c_uint64()
        check(lib.ll_local_topk(self._h, buf, C.byref(n)), self._h)
        k = int(n.value)
        return (np.array([buf[i].fitness for i in range(k)], dtype=np.float64),
                np.array([buf[i].index for i in range(k)], dtype=np.uint64))

    def elite_buffer_bytes(self) -> int:
        return int(lib.ll_elite_buffer_bytes(self._h))

    def gather_elites(self, elite_idx: np.ndarray, device_ptr: int):
        e = np.ascontiguousarray(elite_idx, dtype=np.uint64)
        check(lib.ll_gather_elites(self._h, _p(e, C.c_uint64), len(e), C.c_void_p(device_ptr)), self._h)

    def set_elites(self, device_ptr: int, n: Optional[int] = None):
        check(lib.ll_set_elites(self._h, C.c_void_p(device_ptr), self.top if n is None else n), self._h)

    def mark_elites(self):
        check(lib.ll_mark_elites(self._h), self._h)

    # ---- evaluation
    def set_layouts(self, spad_x1, lpad_x1, directions):
        s = np.ascontiguousarray(spad_x1, dtype=np.int32)
        l = np.ascontiguousarray(lpad_x1, dtype=np.int32)
        d = np.ascontiguousarray(directions, dtype=np.uint64)
        check(lib.ll_set_layouts(self._h, _p(s, C.c_int32), _p(l, C.c_int32), _p(d, C.c_uint64), len(s)), self._h)
        self.n_layouts = len(s)
        check(lib.ll_get_info(self._h, C.byref(self.info)), self._h)  # the episode-kernel variant depends on the layout count

    def evaluate_device(self, max_steps: int):
        check(lib.ll_evaluate_device(self._h, max_steps), self._h)

    def read_results(self, want_arrays: bool = True):
        st = _abi.GenStats()
        if want_arrays:
            f = np.empty(self.pop_local), np.empty(self.pop_local), np.empty(self.pop_local)
            check(lib.ll_read_results(self._h, _p(f[0], C.c_double), _p(f[1], C.c_double), _p(f[2], C.c_double), C.byref(st)), self._h)
            return f[0], f[1], f[2], st
        check(lib.ll_read_results(self._h, None, None, None, C.byref(st)), self._h)
        return None, None, None, st

    def evaluate(self, spad_x1, lpad_x1, directions, max_steps: int, out=None):
        """train_generation: host layouts in, host fitness / lr / steps out (one ABI call)."""
        s = np.ascontiguousarray(spad_x1, dtype=np.int32)
        l = np.ascontiguousarray(lpad_x1, dtype=np.int32)
        d = np.ascontiguousarray(directions, dtype=np.uint64)
        f = out if out is not None else (np.empty(self.pop_local), np.empty(self.pop_local), np.empty(self.pop_local))
        st = _abi.GenStats()
        check(lib.ll_evaluate(self._h, _p(s, C.c_int32), _p(l, C.c_int32), _p(d, C.c_uint64), len(s), max_steps,
                              _p(f[0], C.c_double), _p(f[1], C.c_double), _p(f[2], C.c_double), C.byref(st)), self._h)
        self.n_layouts = len(s)
        check(lib.ll_get_info(self._h, C.byref(self.info)), self._h)
        return f[0], f[1], f[2], st

    def read_episodes(self):
        n = self.pop_local * self.n_layouts
        meta = np.empty(n, dtype=np.int32)
        fit = np.empty(n, dtype=np.float64)
        check(lib.ll_read_episodes(self._h, _p(meta, C.c_int32), _p(fit, C.c_double)), self._h)
        meta = meta.reshape(self.pop_local, self.n_layouts)
        return meta & 0x0FFFFFFF, (meta >> 28) & 7, fit.reshape(self.pop_local, self.n_layouts)

    def step_generation(self, max_steps: int, elitism: bool = True, want_stats: bool = True):
        st = _abi.GenStats()
        check(lib.ll_step_generation(self._h, max_steps, int(elitism), C.byref(st) if want_stats else None), self._h)
        return st

    def trace_episode(self, member: int, spad_x1: int, lpad_x1: int, direction: int, max_steps: int, cap: Optional[int] = None):
        cap = max_steps if cap is None else cap
        rec = (_abi.StepRec * max(cap, 1))()
        res = _abi.EpisodeResult()
        check(lib.ll_trace_episode(self._h, member, spad_x1, lpad_x1, direction, max_steps, rec, cap, C.byref(res)), self._h)
        n = min(cap, res.steps)
        arr = np.frombuffer(rec, dtype=np.float64).reshape(-1, 11)[:n].copy()
        return arr, res

    def feedforward_index(self, state, member: int = 0) -> int:
        s = np.ascontiguousarray(state, dtype=self.dtype)
        a = C.c_uint64()
        check(lib.ll_feedforward_index(self._h, s.ctypes.data_as(C.c_void_p), member, C.byref(a)), self._h)
        return int(a.value)

    # ---- parameters (canonical order: per layer W[out][in] row-major, then b[out])
    def get_members(self, first: int, count: int) -> np.ndarray:
        out = np.empty((count, self.P), dtype=self.dtype)
        check(lib.ll_get_members(self._h, first, count, out.ctypes.data_as(C.c_void_p)), self._h)
        return out

    def set_members(self, first: int, params: np.ndarray):
        p = np.ascontiguousarray(params, dtype=self.dtype).reshape(-1, self.P)
        check(lib.ll_set_members(self._h, first, p.shape[0], p.ctypes.data_as(C.c_void_p)), self._h)

    def serialize(self, path: str, name: str = "lunar_lander", digits: int = 6):
        check(lib.ll_serialize(self._h, path.encode(), name.encode(), digits), self._h)

    def deserialize(self, path: str):
        check(lib.ll_deserialize(self._h, path.encode()), self._h)

    # ---- measurement
    def profile(self, on: bool):
        check(lib.ll_profile(self._h, int(on)), self._h)

    def step_counter(self, reset: bool = False):
        t, n = C.c_double(), C.c_uint64()
        check(lib.ll_step_counter(self._h, C.byref(t), C.byref(n), int(reset)), self._h)
        return float(t.value), int(n.value)

    def kernel_times(self, reset: bool = False) -> _abi.KernelTimes:
        t = _abi.KernelTimes()
        check(lib.ll_kernel_times_get(self._h, C.byref(t), int(reset)), self._h)
        return t


def peek_weight_file(path: str):
    sizes = (C.c_uint64 * 32)()
    n = C.c_int32()
    pop, top, ep = C.c_uint64(), C.c_uint64(), C.c_uint64()
    act, strat = C.c_int32(), C.c_int32()
    ratio = C.c_double()
    name = C.create_string_buffer(256)
    check(lib.ll_peek_weight_file(path.encode(), sizes, 32, C.byref(n), C.byref(pop), C.byref(top), C.byref(ep), C.byref(act),
                                  C.byref(strat), C.byref(ratio), name, 256))
    return dict(sizes=[int(sizes[i]) for i in range(n.value)], population=int(pop.value), top=int(top.value),
                epochs=int(ep.value), act=int(act.value), strategy=int(strat.value), ratio=float(ratio.value),
                name=name.value.decode())


def generate_random_pad_positions(screen_width: float, spad_width: int, lpad_width: int, seed: Optional[int] = None,
                                  force_left_to_right: bool = False, force_right_to_left: bool = False) -> Tuple[int, int]:
    """src/game_utils.h:22-24 / the lunar_lander_cpp binding of the same name (src/lunar_lander_cpp.cpp:98-106)."""
    s, l = C.c_int32(), C.c_int32()
    check(lib.ll_generate_random_pad_positions(screen_width, spad_width, lpad_width, int(seed is not None), 0 if seed is None else seed,
                                               int(force_left_to_right), int(force_right_to_left), C.byref(s), C.byref(l)))
    return int(s.value), int(l.value)


def make_layouts(cfg: Config, base_seed: int, n: int, directions: Optional[np.ndarray] = None):
    """Layout set exactly as train() builds it (src/nn_engine_trainer.cpp:283-309): seeds base+i, then balance_layouts.
    `directions` is in/out state across calls (stale after an early return)."""
    width = cfg.getDouble("DisplayCfg.width")
    sw, lw = cfg.getDouble("GameCfg.spad_width"), cfg.getDouble("GameCfg.lpad_width")
    spad = np.zeros(n, dtype=np.int32)
    lpad = np.zeros(n, dtype=np.int32)
    for i in range(n):
        seed = (base_seed + i) & 0xFFFFFFFF
        seed = seed - (1 << 32) if seed >= (1 << 31) else seed  # static_cast<int>(unsigned)
        spad[i], lpad[i] = generate_random_pad_positions(width, int(sw), int(lw), seed)
    sc = spad.astype(np.float64) + sw / 2.0
    lc = lpad.astype(np.float64) + lw / 2.0
    d = np.zeros(n, dtype=np.uint64) if directions is None else np.ascontiguousarray(directions, dtype=np.uint64)
    check(lib.ll_balance_layouts(_p(spad, C.c_int32), _p(lpad, C.c_int32), _p(sc, C.c_double), _p(lc, C.c_double), _p(d, C.c_uint64), n,
                                 cfg.getDouble("NNTrainingCfg.left_right_ratio"), width, sw, lw))
    return spad, lpad, d, sc, lc
