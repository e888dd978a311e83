"""ORACLE / TEST INFRASTRUCTURE — not part of the product.

ctypes access to the two CPU checkers:
  * ``port``  oracle/_build/libll_oracle.so — our CPU restatement (oracle/ll_oracle.cpp), rebuilt anywhere by ``make -C oracle port``
  * ``ref``   oracle/_ref/libll_ref.so      — the reference's own sources compiled unchanged (oracle/ref_harness.cpp); built only
              where /root/reference exists, travels to the GPU box as a prebuilt file
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_PATH = os.path.join(HERE, "_build", "libll_oracle.so")
REF_PATH = os.path.join(HERE, "_ref", "libll_ref.so")
REF_FAST_PATH = os.path.join(HERE, "_ref", "libll_ref_fast.so")


def build(ref: bool = True) -> None:
    subprocess.run(["make", "-s", "-C", HERE, "port"], check=True)
    if ref and os.path.isdir("/root/reference/src"):
        subprocess.run(["make", "-s", "-C", HERE, "ref"], check=True)


_port = None
_ref = {}


def port() -> C.CDLL:
    global _port
    if _port is None:
        if not os.path.exists(PORT_PATH):
            build(ref=False)
        _port = C.CDLL(PORT_PATH)
        _port.orc_tanh.restype = C.c_double
        _port.orc_tanh.argtypes = [C.c_double]
        _port.orc_tanh_n.restype = None
        _port.orc_tanh_n.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        _port.orc_sigmoid.restype = C.c_double
        _port.orc_sigmoid.argtypes = [C.c_double]
        _port.orc_parent_pick.restype = C.c_uint32
    return _port


def ref(fast: bool = False):
    """The compiled reference, or None when it is not available (e.g. never built in this checkout)."""
    path = REF_FAST_PATH if fast else REF_PATH
    if path not in _ref:
        _ref[path] = C.CDLL(path) if os.path.exists(path) else None
        if _ref[path] is not None:
            _ref[path].ref_last_error.restype = C.c_char_p
    return _ref[path]


def _ptr(a, ty):
    return a.ctypes.data_as(C.POINTER(ty))


def _sizes(sizes):
    return (C.c_uint64 * len(sizes))(*sizes)


def evaluate(env_cfg, sizes, act, params, spad, lpad, dirs, max_steps, precision=64, threads=1):
    """Port: train_generation over `params` [pop][P] (canonical order).  Returns dict of arrays."""
    dt = np.float64 if precision == 64 else np.float32
    p = np.ascontiguousarray(params, dtype=dt)
    pop = p.shape[0]
    spad = np.ascontiguousarray(spad, dtype=np.int32)
    lpad = np.ascontiguousarray(lpad, dtype=np.int32)
    dirs = np.ascontiguousarray(dirs, dtype=np.uint64)
    L = len(spad)
    fit, lr, steps = np.empty(pop), np.empty(pop), np.empty(pop)
    meta = np.empty((pop, L), dtype=np.int32)
    efit = np.empty((pop, L))
    sec = C.c_double()
    fn = port().orc_evaluate_f64 if precision == 64 else port().orc_evaluate_f32
    rc = fn(C.byref(env_cfg), _sizes(sizes), len(sizes), act, C.c_uint64(pop), p.ctypes.data_as(C.c_void_p), _ptr(spad, C.c_int32),
            _ptr(lpad, C.c_int32), _ptr(dirs, C.c_uint64), L, max_steps, threads, _ptr(fit, C.c_double), _ptr(lr, C.c_double),
            _ptr(steps, C.c_double), _ptr(meta, C.c_int32), _ptr(efit, C.c_double), C.byref(sec))
    assert rc == 0
    return dict(fitness=fit, lr=lr, steps=steps, ep_steps=meta & 0x0FFFFFFF, ep_flags=(meta >> 28) & 7, ep_fit=efit, seconds=sec.value)


def trace(env_cfg, sizes, act, params, spad, lpad, direction, max_steps, precision=64, fixed_action=-1, cap=None):
    dt = np.float64 if precision == 64 else np.float32
    cap = max_steps if cap is None else cap
    rec = np.zeros((max(cap, 1), 11))
    n, fl = C.c_int(), C.c_int()
    acc, term = C.c_double(), C.c_double()
    bonus = (C.c_double * 2)()
    pp = None if params is None else np.ascontiguousarray(params, dtype=dt).ctypes.data_as(C.c_void_p)
    fn = port().orc_trace_f64 if precision == 64 else port().orc_trace_f32
    rc = fn(C.byref(env_cfg), _sizes(sizes), len(sizes), act, pp, int(spad), int(lpad), C.c_uint64(direction), max_steps, fixed_action,
            _ptr(rec, C.c_double), cap, C.byref(n), C.byref(fl), C.byref(acc), C.byref(term), bonus)
    assert rc == 0
    return dict(rec=rec[:min(cap, n.value)], steps=n.value, flags=fl.value, acc=acc.value, term=term.value, bonus=(bonus[0], bonus[1]))


def create_population(sizes, seed, pop, top, gen, strategy, ratio, elitism, has_elites, sigma_init, sigma_mut, params, precision=64):
    dt = np.float64 if precision == 64 else np.float32
    p = np.ascontiguousarray(params, dtype=dt).copy()
    fn = port().orc_create_population_f64 if precision == 64 else port().orc_create_population_f32
    fn(_sizes(sizes), len(sizes), C.c_uint32(seed), C.c_uint64(pop), C.c_uint64(top), C.c_uint32(gen), strategy, C.c_double(ratio), int(elitism),
       int(has_elites), C.c_double(sigma_init), C.c_double(sigma_mut), p.ctypes.data_as(C.c_void_p))
    return p


def set_activation_impl(libm: bool) -> None:
    """False (default): the deterministic ll_tanh of csrc/ll_math.h; True: libm's std::tanh evaluated in T (SURVEY.md H2 cross-check)."""
    port().orc_set_activation_impl(1 if libm else 0)


def topk(fitness, k):
    f = np.ascontiguousarray(fitness, dtype=np.float64)
    out = np.zeros(k, dtype=np.uint64)
    port().orc_topk(_ptr(f, C.c_double), C.c_uint64(len(f)), C.c_uint64(k), _ptr(out, C.c_uint64))
    return out


def forward(sizes, act, params, state, precision=64):
    dt = np.float64 if precision == 64 else np.float32
    p = np.ascontiguousarray(params, dtype=dt)
    s = np.ascontiguousarray(state, dtype=dt)
    fn = port().orc_forward_f64 if precision == 64 else port().orc_forward_f32
    return fn(_sizes(sizes), len(sizes), act, p.ctypes.data_as(C.c_void_p), s.ctypes.data_as(C.c_void_p))


# ---- the compiled reference (oracle/_ref) ------------------------------------------------------------------------
def _overrides(sizes, act, max_steps, extra=None):
    kv = {"NNConfig.hlayers": ",".join(str(s) for s in sizes[1:-1]), "NNConfig.activation_id": act, "GameCfg.max_steps": max_steps,
          "NNTrainingCfg.multithread": "true"}
    kv.update(extra or {})
    return ";".join(f"{k}={v}" for k, v in kv.items()).encode()


def ref_evaluate(config_path, sizes, act, params, spad, lpad, dirs, max_steps, precision=64, threads=None, fast=False, extra=None):
    """Reference train_generation (src/nn_engine_trainer.cpp:443-539, unmodified) on the given members."""
    lib = ref(fast)
    assert lib is not None, "oracle/_ref is not built"
    dt = np.float64 if precision == 64 else np.float32
    p = np.ascontiguousarray(params, dtype=dt)
    pop = p.shape[0]
    spad = np.ascontiguousarray(spad, dtype=np.int32)
    lpad = np.ascontiguousarray(lpad, dtype=np.int32)
    dirs = np.ascontiguousarray(dirs, dtype=np.uint64)
    fit, lr, steps = np.empty(pop), np.empty(pop), np.empty(pop)
    calls, sec = C.c_uint64(), C.c_double()
    ex = dict(extra or {})
    if threads is not None:
        os.environ["LL_REF_THREADS"] = str(threads)
        if threads == 1:
            ex["NNTrainingCfg.multithread"] = "false"
    fn = lib.ref_evaluate_f64 if precision == 64 else lib.ref_evaluate_f32
    rc = fn(config_path.encode(), _overrides(sizes, act, max_steps, ex), C.c_uint64(pop), p.ctypes.data_as(C.c_void_p), _ptr(spad, C.c_int32),
            _ptr(lpad, C.c_int32), _ptr(dirs, C.c_uint64), len(spad), _ptr(fit, C.c_double), _ptr(lr, C.c_double), _ptr(steps, C.c_double),
            C.byref(calls), C.byref(sec))
    if rc != 0:
        raise RuntimeError(lib.ref_last_error().decode())
    return dict(fitness=fit, lr=lr, steps=steps, ff_calls=calls.value, seconds=sec.value)


def ref_trace(config_path, sizes, act, params, spad, lpad, direction, max_steps, precision=64, fixed_action=-1, cap=None, extra=None):
    lib = ref()
    assert lib is not None, "oracle/_ref is not built"
    dt = np.float64 if precision == 64 else np.float32
    cap = max_steps if cap is None else cap
    rec = np.zeros((max(cap, 1), 11))
    n, fl = C.c_int(), C.c_int()
    acc, term = C.c_double(), C.c_double()
    bonus = (C.c_double * 2)()
    pp = None if params is None else np.ascontiguousarray(params, dtype=dt).ctypes.data_as(C.c_void_p)
    fn = lib.ref_trace_f64 if precision == 64 else lib.ref_trace_f32
    rc = fn(config_path.encode(), _overrides(sizes, act, max_steps, extra), pp, int(spad), int(lpad), C.c_uint64(direction), fixed_action,
            _ptr(rec, C.c_double), cap, C.byref(n), C.byref(fl), C.byref(acc), C.byref(term), bonus)
    if rc != 0:
        raise RuntimeError(lib.ref_last_error().decode())
    return dict(rec=rec[:min(cap, n.value)], steps=n.value, flags=fl.value, acc=acc.value, term=term.value, bonus=(bonus[0], bonus[1]))


def ref_make_layouts(config_path, base_seed, n, directions=None, extra=None):
    lib = ref()
    assert lib is not None
    spad, lpad = np.zeros(n, dtype=np.int32), np.zeros(n, dtype=np.int32)
    d = np.zeros(n, dtype=np.uint64) if directions is None else np.ascontiguousarray(directions, dtype=np.uint64).copy()
    ov = ";".join(f"{k}={v}" for k, v in (extra or {}).items()).encode()
    rc = lib.ref_make_layouts(config_path.encode(), ov, C.c_uint(base_seed), n, _ptr(spad, C.c_int32), _ptr(lpad, C.c_int32), _ptr(d, C.c_uint64))
    if rc != 0:
        raise RuntimeError(lib.ref_last_error().decode())
    return spad, lpad, d


def ref_threads() -> int:
    lib = ref()
    return int(lib.ref_threads()) if lib is not None else 0


def ref_nn_play(config_path, weight_file, seed=None, force_ltr=False, force_rtl=False, fps=60.0, max_frames=6000, extra=None):
    """main.py --mode=nn_play headless, on the reference's own GameLogicCpp<double> (ref_harness.cpp: ref_nn_play)."""
    lib = ref()
    oi = (C.c_int * 6)()
    of = (C.c_double * 4)()
    rec = np.zeros((max_frames, 6))
    ov = ";".join(f"{k}={v}" for k, v in (extra or {}).items())
    lib.ref_nn_play.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int,
                                C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int]
    lib.ref_last_error.restype = C.c_char_p
    rc = lib.ref_nn_play(config_path.encode(), ov.encode(), weight_file.encode(), int(seed is not None), 0 if seed is None else seed,
                         int(force_ltr), int(force_rtl), fps, max_frames, oi, of, rec.ctypes.data_as(C.POINTER(C.c_double)), max_frames)
    if rc:
        raise RuntimeError(lib.ref_last_error().decode())
    n = oi[3]
    return dict(spad_x1=oi[0], lpad_x1=oi[1], frames=oi[2], steps=n, start_frame=oi[4], landed=bool(oi[5] & 1), crashed=bool(oi[5] & 2),
                landed_successfully=bool(oi[5] & 4), game_over=bool(oi[5] & 8), penalty=of[0], x=of[1], y=of[2], fuel=of[3],
                actions=rec[:n, 0].astype(int), states=rec[:n, 1:].copy())
