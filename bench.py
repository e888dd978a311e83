#!/usr/bin/env python
"""bench.py — lander-steps/sec and generations/sec of the neuro-evolution hot path (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c4|c2|c3|c1|c5|c4_srand] [--scaling weak|strong] [--impl reference]

A "step" is one generation over one batch of synthetic input: evaluate every (member, layout) episode of the
population (K1), select the elites (K2/K3) and breed the next population (K4).  One JSON line is printed by rank 0.

Workloads (`--workload`):
  c4  (default) BASELINE.json configs[3]: 32_128_32, 131072 members PER GPU (= the 1M population at 8 GPUs; weak
      scaling), config.txt's 50 layouts, S-elite population (sample-run elites + N(0, 0.05^2), Philox on the device),
      fp32 fast mode — the configuration the metric's target is stated on
  c2  configs[1]: 8_16_8, population 4096, single start position          (working set < L2: L2 is flushed between steps)
  c3  configs[2]: 16_64_32, population 65536 x 8 start positions
  c1  configs[0]: default config.txt (16_16, population 100, 50 layouts), fp64 exact mode
  c4_srand  the c4 topology on an S-rand population (every weight iid N(0,1), redrawn each generation: bimodal 9-step / 5000-step
      episodes, SURVEY.md §8d "stresses divergence"), 16384 members per GPU
At N = 1 the default line also carries `workloads` = {c1, c2, c3, c5, c4_srand}: short runs of the other BASELINE configs with their own
value / roofline / clocks / cpu_baseline, and every line carries `invariance`: a hash of the fitness vector, elite indices and
children of one small seeded generation sharded over the N ranks — it must be the same string at N = 1, 2, 4, 8.
`--scaling strong` fixes the GLOBAL population (c4: 2^20 members) and splits it over the ranks.
Keys beyond the base contract: `roofline` (episode kernel against the FP32/FP64 FMA peak measured on this GPU, see
DESIGN.md §Measurement), `roofline_ga` (breed kernel against the measured HBM copy bandwidth), `cpu_baseline`
(the reference's own train_generation, oracle/_ref, on this box's host cores over a bounded sample), `e2e`
(the same generations driven through the reference-facing calls with HOST buffers).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

TOPOLOGIES = {
    "16_16": [5, 16, 16, 4], "8_16_8": [5, 8, 16, 8, 4], "16_64_32": [5, 16, 64, 32, 4],
    "32_128_32": [5, 32, 128, 32, 4], "8_16_32_16_8": [5, 8, 16, 32, 16, 8, 4],
}
# name -> (topology, members per GPU, layouts, precision, mode, scaling, description)
WORKLOADS = {
    "c4": ("32_128_32", 131072, 50, 32, 1, "weak",
           "configs[3]: 32_128_32, 131072 members/GPU (1M at 8 GPUs) x 50 layouts, S-elite, fp32 fast"),
    "c3": ("16_64_32", 65536, 8, 32, 1, "weak", "configs[2]: 16_64_32, population 65536/GPU x 8 start positions, S-elite, fp32 fast"),
    "c2": ("8_16_8", 4096, 1, 32, 1, "weak", "configs[1]: 8_16_8, population 4096/GPU, single start position, S-elite, fp32 fast"),
    "c1": ("16_16", 100, 50, 64, 0, "weak", "configs[0]: config.txt defaults, 16_16, population 100/GPU x 50 layouts, S-elite, fp64 exact"),
    "c5": ("8_16_32_16_8", 100, 50, 32, 1, "weak",
           "configs[4] population: 8_16_32_16_8 deep MLP, config.txt population 100/GPU x 50 layouts, S-elite, fp32 fast (what main_train_f trains)"),
    "c4_srand": ("32_128_32", 16384, 50, 32, 1, "weak",
                 "configs[3] topology on S-rand: 32_128_32, 16384 members/GPU x 50 layouts, every weight iid N(0,1) redrawn each generation, fp32 fast"),
}
STRONG_GPUS = 8  # --scaling strong: the global population is the weak-scaling population of 8 GPUs (c4: 2^20 members)
L2_BYTES = 126 << 20
SIGMA_MUT = 0.05


def flop_per_step(sizes):
    mac = sum(sizes[i] * sizes[i + 1] for i in range(len(sizes) - 1))
    return 2 * mac + sum(sizes[1:])


def load_elites(topo):
    z = np.load(os.path.join(ROOT, "tests", "golden", f"elites_{topo}.npz"))
    return z["params"]


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.thread = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        self.thread.join(timeout=5)
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])), mx.append(float(r[1])), pw.append(float(r[2]))
            except ValueError:
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)), "samples": len(sm),
                "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------------------------
# CPU side: the reference's own train_generation (src/nn_engine_trainer.cpp:443-539, compiled unchanged into
# oracle/_ref by oracle/Makefile) on this box's host cores.  Only this leg and --impl reference touch oracle/.
def s_elite_sample(topo, n, seed):
    e = load_elites(topo)
    rng = np.random.default_rng(seed)
    return e[np.arange(n) % e.shape[0]] + rng.normal(scale=SIGMA_MUT, size=(n, e.shape[1]))


def cpu_layouts(n_layouts):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle as po
    if po.ref() is not None:
        return po.ref_make_layouts(os.path.join(ROOT, "config.txt"), 5247, n_layouts)
    raise RuntimeError("oracle/_ref is not built (it is compiled where /root/reference exists and travels prebuilt)")


def cpu_reference_pass(topo, n_layouts, precision, members, layouts, max_steps=5000, threads=None):
    """One pass of the reference CPU path over `members`; returns (lander steps, seconds, threads).  threads=1 is the reference
    with NNTrainingCfg.multithread = false (src/nn_engine_trainer.cpp:533-535)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle as po
    sizes = TOPOLOGIES[topo]
    spad, lpad, d = layouts
    r = po.ref_evaluate(os.path.join(ROOT, "config.txt"), sizes, 1, members, spad, lpad, d, max_steps, precision, threads=threads, fast=True)
    return float(r["steps"].sum()), float(r["seconds"]), (threads or po.ref_threads())


def cpu_calibrate(topo, n_layouts, precision, target_s, seed=99, threads=None, srand=False):
    """Pick a member count whose reference pass takes about `target_s` seconds on `threads` (default: all) host threads."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import pyoracle as po
    nthreads = threads or max(1, po.ref_threads())
    layouts = cpu_layouts(n_layouts)
    n0 = nthreads
    while True:  # grow the probe until fixed costs (thread pool, simulator construction) stop dominating it
        steps0, sec0, _ = cpu_reference_pass(topo, n_layouts, precision, cpu_sample(topo, n0, seed, srand), layouts, threads=threads)
        if sec0 >= 0.4 or n0 >= (1 << 15):
            break
        n0 *= 4
    per_member = max(sec0, 1e-4) / n0
    n = int(max(nthreads, min(1 << 16, round(target_s / per_member / nthreads) * nthreads)))
    return n, layouts, nthreads


def cpu_sample(topo, n, seed, srand=False):
    if not srand:
        return s_elite_sample(topo, n, seed)
    sizes = TOPOLOGIES[topo]
    p = sum(sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(len(sizes) - 1))
    return np.random.default_rng(seed).normal(size=(n, p))


def cpu_baseline(topo, n_layouts, precision, target_s=15.0, threads=None, srand=False):
    try:
        n, layouts, nthreads = cpu_calibrate(topo, n_layouts, precision, target_s, threads=threads, srand=srand)
        steps, sec, _ = cpu_reference_pass(topo, n_layouts, precision, cpu_sample(topo, n, 100, srand), layouts, threads=threads)
    finally:
        os.environ.pop("LL_REF_THREADS", None)
    kind = "S-rand" if srand else "S-elite"
    return {"value": steps / sec, "unit": "lander-steps/s", "cores": nthreads, "kind": "reference",
            "sample": f"{n} {kind} members x {n_layouts} layouts of the same workload ({int(steps)} lander steps in {sec:.2f} s), "
                      f"reference train_generation compiled unchanged (oracle/_ref, g++ -O3 -mavx2 -ffp-contract=off), "
                      f"T={'double' if precision == 64 else 'float'}, policy network from the ma-libs stand-in (oracle/shim)"
                      + (", NNTrainingCfg.multithread = false" if threads == 1 else "")}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path, all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    topo, per_gpu, n_layouts, precision, mode, scaling, desc = WORKLOADS[args.workload]
    n_layouts = args.layouts or n_layouts
    try:
        n, layouts, threads = cpu_calibrate(topo, n_layouts, precision, args.ref_step_seconds, srand=args.workload == "c4_srand")
    except Exception as e:  # the prebuilt oracle/_ref did not travel
        print(json.dumps({"impl": "reference", "unavailable": str(e).splitlines()[0][:200]}))
        return
    tot_steps, tot_sec = 0.0, 0.0
    for it in range(args.warmup + args.steps):
        members = cpu_sample(topo, n, 1000 + it, args.workload == "c4_srand")
        steps, sec, _ = cpu_reference_pass(topo, n_layouts, precision, members, layouts)
        if it >= args.warmup:
            tot_steps += steps
            tot_sec += sec
    value = tot_steps / tot_sec
    sample = (f"each step = reference train_generation over {n} {'S-rand' if args.workload == 'c4_srand' else 'S-elite'} members x {n_layouts} layouts "
              f"({int(tot_steps / args.steps)} lander steps/step), {threads} host threads")
    line = {
        "impl": "reference", "metric": "lander-steps/sec", "value": value, "unit": "lander-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_sec / args.steps, "higher_is_better": True,
        "scaling": scaling, "vs_baseline": None, "dtype": "f64" if precision == 64 else "f32", "data": "synthetic",
        "config": {"workload": desc, "topology": topo, "layouts": n_layouts, "max_steps": 5000, "members_per_step": n},
        "cpu_baseline": {"value": value, "unit": "lander-steps/s", "cores": threads, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "lander-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "generations_per_sec_at_full_population": value / (tot_steps / args.steps / n * per_gpu * max(1, args.gpus)),
    }
    print(json.dumps(line))




# ------------------------------------------------------------------------------------------------------------------
def measured_hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
        except Exception:
            pass
    return 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"


def ncu_traffic(key, members):
    """DRAM bytes per launch from the committed ncu --set full capture (profiles/traffic.json: bytes per member of the captured
    launch x the members of this launch — both kernels touch every member row exactly once), if there is one."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(p):
        try:
            e = json.load(open(p)).get(key)
            return float(e["bytes_per_member"]) * members if e else None
        except Exception:
            return None
    return None


# HMMA work the legacy tensor-core episode kernels really issue per lander step (ll_episode_mma.cu): per layer
# ceil(in/16) * ceil(out/8) tiles x 3 split products x 16*8*16 MACs, for a 16-slot tile
# (ll_episode_mmat.cu, slots on the N side: ceil(in/16) * ceil(out/16) tiles for an 8-slot tile)
def hmma_flop_per_step(sizes, transposed=False):
    if transposed:
        # the first layer (<= 8 inputs) uses m16n8k8: half a tile of work
        tiles = sum(((sizes[i] + 15) // 16) * ((sizes[i + 1] + 15) // 16) * (0.5 if i == 0 and sizes[0] <= 8 else 1.0)
                    for i in range(len(sizes) - 1))
        return tiles * 3 * 2 * 16 * 8 * 16 / 8.0
    tiles = sum(((sizes[i] + 15) // 16) * ((sizes[i + 1] + 7) // 8) for i in range(len(sizes) - 1))
    return tiles * 3 * 2 * 16 * 8 * 16 / 16.0


def tc05_flop_per_step(sizes):
    """tcgen05.mma flop per live row (= lander step) of ll_episode_tc05_kernel: per layer (K blocks of 16) x 3 split products (the first
    layer packs its three products into 2 instructions) x N padded (hidden layers to 16, the output layer to 8) x K = 16 x 2."""
    n_layers, flop = len(sizes) - 1, 0
    for l in range(n_layers):
        npad = 8 if l == n_layers - 1 else (sizes[l + 1] + 15) // 16 * 16
        instr = 2 if l == 0 else 3 * ((sizes[l] + 15) // 16)
        flop += instr * npad * 16 * 2
    return flop


def measured_bf16_peak(key="bf16_tflops"):
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))[key])
    except Exception:
        return None


class Peaks:
    """Roofline denominators measured on THIS device (rank-local), each once per run: FMA chains (ll_microbench_fma), mean over a
    0.25 s burst and over a 3 s loop, with the SM clock the same thread saw (cycles / %globaltimer ns)."""

    def __init__(self, abi, device):
        self.abi, self.device, self.cache = abi, device, {}

    def get(self, code, seconds):
        import ctypes as C
        k = (code, seconds)
        if k not in self.cache:
            v, mhz = C.c_double(), C.c_double()
            self.abi.check(self.abi.lib.ll_microbench_fma(self.device, code, seconds, C.byref(v), C.byref(mhz)))
            self.cache[k] = (v.value, mhz.value)
        return self.cache[k]


def measure(ctx, workload, steps, warmup, min_timed_s=1.5, e2e_steps=None, cpu_seconds=15.0, want_cpu=True):
    """One workload on this rank's GPU (all ranks call it together).  Returns the result dict on rank 0, None elsewhere."""
    import torch
    import torch.distributed as dist
    ll, abi, rank, world, local_rank, stream, peaks, args = (ctx["ll"], ctx["abi"], ctx["rank"], ctx["world"], ctx["local_rank"],
                                                             ctx["stream"], ctx["peaks"], ctx["args"])
    from lunar_lander_b200.sharded import ShardedGeneration

    topo, per_gpu, n_layouts, precision, mode, scaling, desc = WORKLOADS[workload]
    srand = workload == "c4_srand"
    if workload == args.workload:
        per_gpu = args.pop_per_gpu or per_gpu
        n_layouts = args.layouts or n_layouts
        if args.precision:
            precision, mode = args.precision, (0 if args.precision == 64 or args.exact else 1)
        elif args.exact:
            mode = 0
        if args.scaling == "strong":
            scaling, per_gpu = "strong", (per_gpu * STRONG_GPUS) // world
    sizes = TOPOLOGIES[topo]
    top = 10
    cfg = ll.Config(os.path.join(ROOT, "config.txt"), {"NNConfig.hlayers": ",".join(str(s) for s in sizes[1:-1])})
    max_steps = cfg.getInt("GameCfg.max_steps")
    spad, lpad, dirs, _, _ = ll.make_layouts(cfg, 5247, n_layouts)

    pcode = 64 if precision == 64 else 32
    fma_burst, mhz_b = peaks.get(pcode, 0.25)
    fma_sust, mhz_s = peaks.get(pcode, 3.0)
    hmma_peak = peaks.get(197, 0.25)[0] if (precision == 32 and mode == 1) else 0.0

    eng = ll.Engine(cfg, sizes=sizes, population=per_gpu * world, top=top, precision=precision, mode=mode, rank=rank, world=world,
                    device=local_rank, seed=5247, sigma_mut=SIGMA_MUT)
    eng.set_stream(stream.cuda_stream)
    sg = ShardedGeneration(eng)  # world > 1: hands rank 0's communicator id to the others, ll_comm_init

    if srand:
        # S-rand: K0 draws every weight iid N(0,1); with an injection ratio of 1 every timed generation redraws all children
        eng.set_population_strategy(2, 1.0)
        eng.create_population(True)
    else:
        # S-elite, generated on the device: parents = the sample run's elites (rank 0 owns slots [0, top); ll_select with given
        # indices gathers them from their owner and all-reduces), children = parent + N(0, 0.05^2) by Philox (K4); the timed
        # generations then use config.txt's strategy
        if rank == 0:
            eng.set_members(0, load_elites(topo)[:top])
        eng.select(np.arange(top, dtype=np.uint64))
        eng.set_population_strategy(1, 0.0)
        eng.create_population(True)
        eng.set_population_strategy(2, cfg.getDouble("NNConfig.random_injection_ratio"))
    eng.set_layouts(spad, lpad, dirs)
    eng.synchronize()

    row_bytes = int(eng.info.params_padded) * (8 if precision == 64 else 4)
    working_set = eng.pop_local * row_bytes
    flush = working_set < 2 * L2_BYTES
    flush_buf = ctx.setdefault("flush_buf", torch.empty(512 << 20, dtype=torch.uint8, device="cuda")) if flush else None

    def generation():
        sg.step(max_steps, True)  # one C-ABI call: ll_step_generation (K1, K2, [NCCL all-gather, merge,] K3, [NCCL all-reduce,] K4)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    t_w0 = time.perf_counter()
    for _ in range(warmup):
        generation()
    barrier()
    warm_s = (time.perf_counter() - t_w0) / max(1, warmup)
    if workload != args.workload or args.auto_steps:  # short workloads: enough steps for >= 10 clock samples
        steps = int(max(steps, min(20000, np.ceil(min_timed_s / max(warm_s, 1e-5)))))
        if world > 1:
            t = torch.tensor([steps], dtype=torch.int64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            steps = int(t[0])
    eng.step_counter(reset=True)
    eng.kernel_times(reset=True)
    eng.profile(True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pairs = []
    barrier()
    t_wall0 = time.perf_counter()
    if flush:  # the L2 flush sits between timed steps, outside the per-step event pairs
        for k in range(steps):
            flush_buf.fill_(k & 255)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            generation()
            b.record(stream)
            pairs.append((a, b))
    else:
        ev0.record(stream)
        for k in range(steps):
            generation()
        ev1.record(stream)
    barrier()
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop() if rank == 0 else None
    ms_total = sum(a.elapsed_time(b) for a, b in pairs) if flush else ev0.elapsed_time(ev1)
    steps_local, n_eval = eng.step_counter(reset=True)
    kt = eng.kernel_times(reset=True)
    eng.profile(False)
    t = torch.tensor([ms_total, kt.episode_ms], dtype=torch.float64, device="cuda")
    s = torch.tensor([steps_local], dtype=torch.float64, device="cuda")
    k1_all = torch.zeros(world, dtype=torch.float64, device="cuda")
    k1_all[rank] = kt.episode_ms / max(1, steps)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
        dist.all_reduce(k1_all, op=dist.ReduceOp.SUM)
    ms_total_max, episode_ms_max = float(t[0]), float(t[1])
    steps_global = float(s[0])
    launches = int(kt.episode_launches + kt.select_launches + kt.gather_launches + kt.mutate_launches + kt.other_launches)

    # ---- e2e: the same generations driven the way the reference's train() drives them (src/nn_engine_trainer.cpp:373-407)
    # with HOST buffers: layouts go in from host memory, fitness / lr / steps come back to host vectors, the host sorts the
    # indices (std::sort in the reference) and hands the order to UpdateWeightsAndBiases.  Copies are inside the timed region.
    # With several ranks the sort is the device top-k + NCCL merge behind ll_select(NULL) (a host sort would need the whole vector).
    # The sort: what the shipped trainer does at this size (host/nn_engine_trainer.cpp, GpuCfg.select = auto): the reference's host
    # std::sort up to 65 536 members, the device top-k (ll_select(NULL), elite indices back to the host) above.
    host_sort = eng.pop_local * world <= 65536
    n_e2e = e2e_steps if e2e_steps is not None else steps
    out = (np.empty(eng.pop_local), np.empty(eng.pop_local), np.empty(eng.pop_local))
    h2d = d2h = 0
    e2e_lander_steps = 0.0
    barrier()
    t0 = time.perf_counter()
    for k in range(n_e2e):
        fit, lr, st_steps, st = eng.evaluate(spad, lpad, dirs, max_steps, out=out)
        h2d += n_layouts * 16
        d2h += 3 * 8 * eng.pop_local + 48
        e2e_lander_steps += float(st_steps.sum())
        if world == 1:
            if host_sort:
                order = np.argsort(fit, kind="stable").astype(np.uint64)
                eng.select(order)
                h2d += 8 * top
            else:
                eng.select(None)
                d2h += 8 * top
            eng.create_population(True)
            eng.update_epochs(1)
        else:
            b1 = sg.d2h_bytes
            sg.select_and_breed(True)
            d2h += sg.d2h_bytes - b1
    barrier()
    e2e_sec = time.perf_counter() - t0
    t = torch.tensor([e2e_sec], dtype=torch.float64, device="cuda")
    s = torch.tensor([e2e_lander_steps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(s, op=dist.ReduceOp.SUM)
    e2e_sec, e2e_lander_steps = float(t[0]), float(s[0])

    res = None
    if rank == 0:
        kernel = eng.info.kernel.decode()
        fps = flop_per_step(sizes)
        value = steps_global / (ms_total_max * 1e-3)
        achieved = (steps_local * fps) / (episode_ms_max * 1e-3) / 1e12 if episode_ms_max > 0 else 0.0
        long_kernel = episode_ms_max / max(1, steps) > 250.0
        peak = fma_sust if long_kernel else fma_burst
        nominal = (148 * 128 * 2 * 1.965e9 / 1e12) if precision == 32 else None
        hbm_peak, hbm_src = measured_hbm_peak()
        elt = 8 if precision == 64 else 4
        ga_bytes = ((eng.pop_local - top) * int(eng.info.params_padded) * elt + top * int(eng.info.params_padded) * elt) * steps
        ga_gbs = ga_bytes / (kt.mutate_ms * 1e-3) / 1e9 if kt.mutate_ms > 0 else 0.0
        k1 = [float(x) for x in k1_all.cpu().numpy()]
        res = {
            "value": value, "unit": "lander-steps/s", "steps": steps, "warmup": warmup, "ms_per_step": ms_total_max / steps, "scaling": scaling,
            "dtype": "f64" if precision == 64 else "f32",
            "config": {"workload": desc, "topology": topo, "population": per_gpu * world, "population_per_gpu": per_gpu,
                       "members_per_step": per_gpu * world, "layouts": n_layouts, "max_steps": max_steps, "mode": "exact" if mode == 0 else "fast",
                       "strategy": ("every child a fresh N(0,1) individual (injection ratio 1), elitism, top 10" if srand else
                                    "MIXED_WITH_RANDOM_INJECTION 0.35, elitism, top 10, sigma_mut 0.05"),
                       "l2": "L2 flushed (512 MiB write) between timed steps" if flush else
                             f"per-GPU working set {working_set / 1e6:.0f} MB of member rows > 126 MB L2, no flush needed",
                       "kernel": kernel},
            "generations_per_sec": steps / (ms_total_max * 1e-3),
            "lander_steps_per_generation": steps_global / steps,
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "fp64_fma" if precision == 64 else "fp32_fma", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                         "frac": achieved / peak if peak > 0 else None,
                         "frac_nominal": achieved / nominal if nominal else None,
                         "traffic": ncu_traffic(workload, eng.pop_local) if "mma" in kernel or "tc05" in kernel else None,
                         "kernel": kernel, "flop_per_lander_step": fps,
                         "kernel_ms_per_launch": episode_ms_max / max(1, steps),
                         "kernel_ms_per_launch_by_rank": k1,
                         "kernel_share_of_step": episode_ms_max / ms_total_max if ms_total_max > 0 else None,
                         "peak_source": ("FMA microbenchmark on this GPU (ll_microbench_fma), mean of a " +
                                         ("3 s loop (the kernel runs > 250 ms per launch)" if long_kernel else "0.25 s burst") + ", of measured"),
                         "peak_burst": fma_burst, "peak_sustained": fma_sust,
                         "sm_mhz_during_peak": mhz_s if long_kernel else mhz_b,
                         "peak_nominal_1965mhz": nominal},
            "tensor_pipe": ({"note": "explains the FMA-roofline fraction: HMMA flop the legacy tensor-core kernel issues (3 f16-split products, "
                                     "padded tiles, every row of the tile counted whether its episode is live or not) against the measured "
                                     "mma.sync.m16n8k16.f16 rate",
                             "hmma_flop_per_lander_step": hmma_flop_per_step(sizes, "mmat" in kernel),
                             "achieved_live_rows": steps_local * hmma_flop_per_step(sizes, "mmat" in kernel) / (episode_ms_max * 1e-3) / 1e12 if episode_ms_max > 0 else None,
                             "peak": hmma_peak, "unit": "TFLOP/s"} if "_mma" in kernel else
                            {"note": "explains the FMA-roofline fraction: flop the tcgen05 tile kernel issues per LIVE lander step (3 f16-split products "
                                     "on K = 16 blocks, N padded to 8 / 16; M = 64 tiles whose dead rows are not counted; the straggler pass's steps "
                                     "are counted at the same rate although they run on mma.sync) against the measured dense bf16 rate of "
                                     "MEASURED_PEAKS.json - an M = 64 tile can use at most half of that",
                             "mma_flop_per_lander_step": tc05_flop_per_step(sizes),
                             "achieved_live_rows": steps_local * tc05_flop_per_step(sizes) / (episode_ms_max * 1e-3) / 1e12 if episode_ms_max > 0 else None,
                             "peak": measured_bf16_peak(), "peak_sustained": measured_bf16_peak("bf16_tflops_sustained"), "unit": "TFLOP/s"}
                            if "_tc05_" in kernel else None),
            "roofline_ga": {"bound": "hbm", "achieved": ga_gbs, "peak": hbm_peak, "unit": "GB/s",
                            "frac": ga_gbs / hbm_peak if hbm_peak else None, "traffic": ncu_traffic("k4_" + topo, eng.pop_local - top),
                            "kernel": ("ll_create_population_f32_kernel<fast: Philox4x32-7 + SFU Box-Muller>" if mode == 1 and precision == 32 else
                                       "ll_create_population_f32_kernel<host-reproducible normals>" if precision == 32 else "ll_create_population_kernel<double>"),
                            "kernel_ms_per_launch": kt.mutate_ms / max(1, steps), "launches_per_generation": kt.mutate_launches / max(1, steps),
                            "peak_source": hbm_src},
            "kernel_ms": {"episode": kt.episode_ms, "select": kt.select_ms, "gather": kt.gather_ms, "breed": kt.mutate_ms,
                          "wall_s_timed_region": t_wall},
            "e2e": {"value": e2e_lander_steps / e2e_sec if e2e_sec > 0 else 0.0, "unit": "lander-steps/s",
                    "h2d_bytes_per_step": h2d / max(1, n_e2e), "d2h_bytes_per_step": d2h / max(1, n_e2e), "steps": n_e2e,
                    "generations_per_sec": n_e2e / e2e_sec if e2e_sec > 0 else 0.0,
                    "api": (("ll_evaluate(host layouts -> host fitness/lr/steps) + host argsort + ll_select(sorted) + ll_create_population" if host_sort else
                             "ll_evaluate(host layouts -> host fitness/lr/steps) + ll_select(NULL) [device top-k, elite indices to the host; GpuCfg.select = "
                             "auto above 65 536 members] + ll_create_population") if world == 1 else
                            "per rank: ll_evaluate(host layouts -> host fitness/lr/steps of the shard) + ll_select(NULL) [device top-k, NCCL all-gather, "
                            "device merge, row gather, NCCL all-reduce; elite indices to the host] + ll_create_population")},
        }
        if want_cpu and world == 1:
            try:
                res["cpu_baseline"] = cpu_baseline(topo, n_layouts, precision, cpu_seconds, srand=srand)
            except Exception as e:
                res["cpu_baseline"] = {"value": None, "unit": "lander-steps/s", "cores": 0, "kind": "reference",
                                       "sample": f"unavailable: {str(e).splitlines()[0][:160]}"}
    eng.close()
    return res


def invariance(ctx):
    """SURVEY.md §4 "multi-GPU invariance 1/2/4/8" on real hardware, untimed: one small seeded generation (16_64_32, 4096 members x 8
    layouts, fp32 EXACT mode: sample-run elites bred once by config.txt's GA) sharded over the N ranks.  sha256 over the merged
    fitness vector, the global elite indices and the next generation's children — the same string for every N."""
    import hashlib
    import torch
    import torch.distributed as dist
    ll, rank, world, local_rank, stream = ctx["ll"], ctx["rank"], ctx["world"], ctx["local_rank"], ctx["stream"]
    from lunar_lander_b200.sharded import ShardedGeneration
    topo, pop, n_layouts, top = "16_64_32", 4096, 8, 10
    sizes = TOPOLOGIES[topo]
    cfg = ll.Config(os.path.join(ROOT, "config.txt"), {"NNConfig.hlayers": ",".join(str(s) for s in sizes[1:-1])})
    spad, lpad, dirs, _, _ = ll.make_layouts(cfg, 5247, n_layouts)
    eng = ll.Engine(cfg, sizes=sizes, population=pop, top=top, precision=32, mode=0, rank=rank, world=world, device=local_rank, seed=5247)
    eng.set_stream(stream.cuda_stream)
    sg = ShardedGeneration(eng)
    if rank == 0:
        eng.set_members(0, load_elites(topo)[:top])
    eng.select(np.arange(top, dtype=np.uint64))
    eng.set_population_strategy(2, 0.35)
    eng.create_population(True)
    fit, lr, steps, _ = eng.evaluate(spad, lpad, dirs, 5000)
    elite_idx = sg.select_and_breed(True)
    children = eng.get_members(eng.first_global, eng.pop_local)

    def merged(a):
        a = np.ascontiguousarray(a)
        if world == 1:
            return a
        mine = torch.from_numpy(a.view(np.uint8).reshape(-1)).cuda()
        every = torch.empty(world * mine.numel(), dtype=torch.uint8, device="cuda")
        dist.all_gather_into_tensor(every, mine)
        return every.cpu().numpy()

    parts = [merged(fit), merged(steps), np.ascontiguousarray(elite_idx), merged(children)]
    eng.close()
    h = hashlib.sha256()
    for p in parts:
        h.update(p.tobytes())
    return {"sha256": h.hexdigest(), "what": f"{topo}, {pop} members x {n_layouts} layouts, fp32 exact mode, sharded over {world} rank(s): fitness + "
                                            "steps vectors, elite indices, children of the next generation",
            "elites": [int(i) for i in elite_idx], "lander_steps": float(np.frombuffer(parts[1].tobytes(), dtype=np.float64).sum())}


def run_gpu(args):
    import torch
    import torch.distributed as dist

    import lunar_lander_b200 as ll
    from lunar_lander_b200 import _abi

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — this framework has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    if args.gpus != world and rank == 0:
        print(f"bench.py: --gpus {args.gpus} but WORLD_SIZE={world}; using {world}", file=sys.stderr)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    ctx = {"ll": ll, "abi": _abi, "rank": rank, "world": world, "local_rank": local_rank, "stream": stream,
           "peaks": Peaks(_abi, local_rank), "args": args}

    inv = None if args.no_invariance else invariance(ctx)
    main = measure(ctx, args.workload, args.steps, args.warmup, e2e_steps=args.e2e_steps, cpu_seconds=args.cpu_seconds,
                   want_cpu=not args.no_cpu_baseline)
    extras = {}
    if world == 1 and not args.no_extras and args.workload == "c4":
        # the other BASELINE configs and the S-rand population, each a short run of its own (>= 1.5 s timed, >= 10 clock samples)
        for w in ("c1", "c2", "c3", "c5", "c4_srand"):
            r = measure(ctx, w, 3, 3, cpu_seconds=args.extra_cpu_seconds, want_cpu=not args.no_cpu_baseline)
            for k in ("tensor_pipe", "kernel_ms"):
                r.pop(k, None)
            extras[w] = r
        if not args.no_cpu_baseline:
            try:  # the reference with NNTrainingCfg.multithread = false on its own default configuration
                extras["c1"]["cpu_baseline_1thread"] = cpu_baseline("16_16", 50, 64, args.extra_cpu_seconds, threads=1)
            except Exception as e:
                extras["c1"]["cpu_baseline_1thread"] = {"value": None, "sample": f"unavailable: {str(e).splitlines()[0][:160]}"}
    if rank == 0:
        line = {"metric": "lander-steps/sec", "value": main.pop("value"), "unit": main.pop("unit"), "n_gpus": world, "steps": main.pop("steps"),
                "warmup": main.pop("warmup"), "ms_per_step": main.pop("ms_per_step"), "higher_is_better": True, "scaling": main.pop("scaling"),
                "vs_baseline": None, "dtype": main.pop("dtype"), "data": "synthetic"}
        line.update(main)
        line["invariance"] = inv
        if extras:
            line["workloads"] = extras
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c4", choices=sorted(WORKLOADS))
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--pop-per-gpu", type=int, default=0)
    ap.add_argument("--layouts", type=int, default=0)
    ap.add_argument("--precision", type=int, default=0, choices=[0, 32, 64])
    ap.add_argument("--exact", action="store_true", help="LL_MODE_EXACT (bit-exact operation order) instead of the fast fp32 mode")
    ap.add_argument("--auto-steps", action="store_true", help="raise --steps until the timed region lasts >= 1.5 s (short workloads)")
    ap.add_argument("--e2e-steps", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the `workloads` runs (c1, c2, c3, c5, c4_srand) of the default line")
    ap.add_argument("--no-invariance", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--extra-cpu-seconds", type=float, default=3.0)
    ap.add_argument("--ref-step-seconds", type=float, default=8.0)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
