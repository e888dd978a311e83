"""Development probe: per-episode step counts of the bench population (after a few GA generations) -> gpurun_out/ep_steps_bench.npy."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import lunar_lander_b200 as ll
import bench

topo, pop, L = "32_128_32", 8192, 50
sizes = bench.TOPOLOGIES[topo]
cfg = ll.Config(os.path.join(ROOT, "config.txt"), {"NNConfig.hlayers": ",".join(str(s) for s in sizes[1:-1])})
spad, lpad, dirs, _, _ = ll.make_layouts(cfg, 5247, L)
eng = ll.Engine(cfg, sizes=sizes, population=pop, top=10, precision=32, mode=1, device=0, seed=5247, sigma_mut=bench.SIGMA_MUT)
eng.set_members(0, bench.load_elites(topo)[:10])
eng.mark_elites()
eng.set_population_strategy(1, 0.0)
eng.create_population(True)
eng.set_population_strategy(2, cfg.getDouble("NNConfig.random_injection_ratio"))
eng.set_layouts(spad, lpad, dirs)
for g in range(4):
    eng.step_generation(5000, True, want_stats=False)
    fit, lr, steps, st = eng.evaluate(spad, lpad, dirs, 5000)
    ep, fl, ef = eng.read_episodes()
    print("generation", g, "steps/member", steps.mean(), "episode mean", ep.mean(), flush=True)
    np.save(os.path.join(ROOT, "gpurun_out", f"ep_steps_bench_g{g}.npy"), ep.astype(np.int32))
np.save(os.path.join(ROOT, "gpurun_out", "layouts_bench.npy"), np.stack([spad, lpad, dirs.astype(np.int32)]))
