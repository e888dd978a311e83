#!/usr/bin/env python
"""One fast-mode evaluation of an S-elite population through the C ABI, for ncu / quick timing:
   LL_FAST_KERNEL=tc05 python scripts/run_episode_kernel.py [topo] [pop] [layouts] [repeats]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lunar_lander_b200 as ll  # noqa: E402

topo = sys.argv[1] if len(sys.argv) > 1 else "32_128_32"
pop = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
n_layouts = int(sys.argv[3]) if len(sys.argv) > 3 else 50
repeats = int(sys.argv[4]) if len(sys.argv) > 4 else 2
sizes = [5] + [int(s) for s in topo.split("_")] + [4]
cfg = ll.Config(os.path.join(ROOT, "config.txt"), {"NNConfig.hlayers": ",".join(str(s) for s in sizes[1:-1])})
spad, lpad, dirs, _, _ = ll.make_layouts(cfg, 5247, n_layouts)
eng = ll.Engine(cfg, sizes=sizes, population=pop, top=10, precision=32, mode=1, device=0, seed=5247, sigma_mut=0.05)
e = np.load(os.path.join(ROOT, "tests", "golden", f"elites_{topo}.npz"))["params"][:10]
eng.set_members(0, e)
eng.mark_elites()
eng.set_population_strategy(2 if os.environ.get("INJECT", "1") == "1" else 1, 0.35)
eng.create_population(True)
eng.set_layouts(spad, lpad, dirs)
for g in range(int(os.environ.get("GENS", "0"))):  # evolve a few generations first: later populations have a wider spread of episode lengths
    eng.step_generation(5000, True, want_stats=False)
for r in range(repeats):
    eng.profile(True)
    eng.kernel_times(reset=True)
    t0 = time.perf_counter()
    eng.evaluate_device(5000)
    _, _, _, st = eng.read_results(want_arrays=False)
    dt = time.perf_counter() - t0
    kt = eng.kernel_times(reset=True)
    print(f"{eng.info.kernel.decode()}: pop {pop} x {n_layouts}: {st.sum_steps:.4e} lander steps, K1 {kt.episode_ms:.2f} ms "
          f"-> {st.sum_steps / (kt.episode_ms * 1e-3):.4e} steps/s (wall {dt * 1e3:.1f} ms)", flush=True)
eng.close()
