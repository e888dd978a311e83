"""One small evaluation through every episode-kernel variant + the GA + the allocation-free top-k + the stepped lander (a quick all-paths check;
it was the compute-sanitizer target of round 1 — the sanitizer is closed on this pool now)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import lunar_lander_b200 as ll  # noqa: E402

cfg_path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "config.txt")


KEYS = ("LL_FAST_KERNEL", "LL_MMAT_TILES", "LL_MMA_CHUNKS", "LL_EXACT_CHUNKS", "LL_EXACT_Q", "LL_EXACT_WARP", "LL_TC05_CUT", "LL_TC05_RESUME")


def run(topo, pop, n_layouts, precision, mode, env=None):
    for k in KEYS:
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


run("16,16", 37, 50, 64, 0)                                   # exact, warp per episode (default at this size)
run("16,16", 37, 50, 64, 0, {"LL_EXACT_WARP": "0"})           # exact, tiled
run("16,16", 37, 50, 64, 0, {"LL_EXACT_CHUNKS": "5"})
run("8,16,8", 21, 13, 32, 0, {"LL_EXACT_Q": "4"})
run("16,64,32", 9, 8, 32, 0, {"LL_EXACT_Q": "16"})
run("32,128,32", 5, 13, 64, 0)
run("32,128,32", 19, 50, 32, 1, {"LL_FAST_KERNEL": "tc05"})                                   # tile kernel + N-side straggler pass
run("32,128,32", 19, 50, 32, 1, {"LL_FAST_KERNEL": "tc05", "LL_TC05_RESUME": "mma", "LL_TC05_CUT": "40"})
run("16,64,32", 45, 33, 32, 1, {"LL_FAST_KERNEL": "tc05", "LL_TC05_CUT": "0"})
run("32,128,32", 19, 50, 32, 1, {"LL_FAST_KERNEL": "mma"})
run("32,128,32", 19, 50, 32, 1, {"LL_FAST_KERNEL": "mma", "LL_MMA_CHUNKS": "1"})
run("16,64,32", 45, 8, 32, 1)
run("16,64,32", 45, 20, 32, 1, {"LL_FAST_KERNEL": "mmat"})
run("8,16,8", 70, 1, 32, 1)                                   # warp per member
run("16,32", 29, 17, 32, 1, {"LL_FAST_KERNEL": "ffma"})
# the allocation-free top-k
import ctypes as C  # noqa: E402
import torch  # noqa: E402
from lunar_lander_b200 import _abi  # noqa: E402
fit = torch.randn(70001, dtype=torch.float64, device="cuda")
out = torch.zeros(2 * 37, dtype=torch.int64, device="cuda")
assert _abi.lib.ll_topk_static(-1, None, C.c_void_p(fit.data_ptr()), fit.numel(), 0, 37, C.c_void_p(out.data_ptr())) == 0
torch.cuda.synchronize()
print("topk_static", out[1::2][:5].tolist())
# the stepped lander through the reference's module name
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lunar_lander_b200", "pymodules"))
import lunar_lander_cpp  # noqa: E402
g = lunar_lander_cpp.GameLogicCppDouble(True, cfg_path)
buf = np.zeros(5)
g.reset(100, 500)
for i in range(50):
    g.update(i % 4, buf)
print("sim", g.x, g.y, g.calculate_step_penalty(1), g.calculate_terminal_penalty(50))
