#!/usr/bin/env python
"""Episode-length structure of the benchmark population (what bounds the tile kernels): for the c4 workload after a few GA
generations, per member the 50 episode lengths -> live-slot fraction of a 64-slot tile that runs until the longest episode ends
(ll_episode_tc05.cu), and what handing the last K live episodes to a straggler pass would recover."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lunar_lander_b200 as ll  # noqa: E402

pop = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
gens = int(sys.argv[2]) if len(sys.argv) > 2 else 4
sizes = [5, 32, 128, 32, 4]
cfg = ll.Config(os.path.join(ROOT, "config.txt"), {"NNConfig.hlayers": "32,128,32"})
spad, lpad, dirs, _, _ = ll.make_layouts(cfg, 5247, 50)
eng = ll.Engine(cfg, sizes=sizes, population=pop, top=10, precision=32, mode=1, device=0, seed=5247, sigma_mut=0.05)
eng.set_members(0, np.load(os.path.join(ROOT, "tests", "golden", "elites_32_128_32.npz"))["params"][:10])
eng.mark_elites()
eng.set_population_strategy(1, 0.0)
eng.create_population(True)
eng.set_population_strategy(2, 0.35)
eng.set_layouts(spad, lpad, dirs)
for g in range(gens):
    eng.evaluate_device(5000)
    ep_steps, flags, _ = eng.read_episodes()
    s = np.sort(ep_steps, axis=1)[:, ::-1].astype(np.float64)  # longest first
    mx = s[:, 0]
    tot = s.sum()
    line = [f"gen {g}: steps/member {tot / pop:.0f}, mean max-len {mx.mean():.0f}, live fraction of a 64-slot tile {tot / (64 * mx.sum()):.3f}"]
    for k in (2, 4, 8, 12, 16):
        cut = s[:, k]  # the tile stops when only k episodes are left: runs for the (k+1)-th longest
        left = (s[:, :k] - cut[:, None]).sum()
        line.append(f"cut@{k}: tile time x{cut.sum() / mx.sum():.3f}, leftover {left / tot:.3f} of the steps")
    print("; ".join(line), flush=True)
    eng.select(None)
    eng.create_population(True)
    eng.update_epochs(1)
eng.close()
