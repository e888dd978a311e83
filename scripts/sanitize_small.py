"""compute-sanitizer target: one small evaluation through every episode-kernel variant + the GA + the stepped lander."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lunar_lander_b200 as ll  # noqa: E402

cfg_path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "config.txt")


def run(topo, pop, n_layouts, precision, mode, env=None):
    for k in ("LL_FAST_KERNEL", "LL_MMAT_MEMBERS", "LL_MMAT_TILES", "LL_MMA_CHUNKS", "LL_EXACT_CHUNKS", "LL_EXACT_Q"):
        os.environ.pop(k, None)
    os.environ.update(env or {})
    cfg = ll.Config(cfg_path, {"NNConfig.hlayers": topo})
    spad, lpad, dirs, _, _ = ll.make_layouts(cfg, 5247, n_layouts)
    eng = ll.Engine(cfg, population=pop, top=4, precision=precision, mode=mode, device=0)
    eng.set_population_strategy(2, 0.35)
    eng.create_population(True)
    fit, lr, steps, st = eng.evaluate(spad, lpad, dirs, 300)
    eng.select(None)
    eng.create_population(True)
    eng.update_epochs(1)
    eng.step_generation(300)
    print(topo, pop, n_layouts, precision, mode, env, eng.info.kernel.decode(), int(steps.sum()))
    eng.close()


run("16,16", 37, 50, 64, 0)
run("16,16", 37, 50, 64, 0, {"LL_EXACT_CHUNKS": "5"})
run("8,16,8", 21, 13, 32, 0, {"LL_EXACT_Q": "4"})
run("32,128,32", 19, 50, 32, 1)
run("32,128,32", 19, 50, 32, 1, {"LL_MMA_CHUNKS": "1"})
run("16,64,32", 45, 8, 32, 1)
run("16,64,32", 45, 8, 32, 1, {"LL_MMAT_MEMBERS": "3"})
run("16,64,32", 45, 20, 32, 1, {"LL_FAST_KERNEL": "mmat"})
run("8,16,32,16,8", 33, 5, 32, 1, {"LL_MMAT_MEMBERS": "4"})
run("16,32", 29, 17, 32, 1, {"LL_FAST_KERNEL": "ffma"})
from lunar_lander_b200 import lunar_lander_cpp  # noqa: E402
g = lunar_lander_cpp.GameLogicCppDouble(True, cfg_path)
buf = np.zeros(5)
g.reset(100, 500)
for i in range(50):
    g.update(i % 4, buf)
print("sim", g.x, g.y, g.calculate_step_penalty(1), g.calculate_terminal_penalty(50))
