"""Development probe: tensor-core episode kernel vs the FFMA kernels vs the CPU oracle on one population."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import lunar_lander_b200 as ll
import pyoracle as po
import util

topo, pop, L = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
members, sizes = util.elite_population(topo, pop, 4, noise=0.02)
members[pop // 2:] += util.random_population(sizes, pop - pop // 2, 8, scale=0.05)
spad, lpad, d = util.layouts(50)
spad, lpad, d = spad[:L], lpad[:L], d[:L]
cfg = util.cfg_for(sizes)
res = {}
for name, envs in (("ffma", {"LL_FAST_KERNEL": "ffma"}), ("mma", {"LL_FAST_KERNEL": "mma"}),
                   ):
    for k in ("LL_FAST_KERNEL", "LL_MMA_TERMS", "LL_MMA_MT"):
        os.environ.pop(k, None)
    os.environ.update(envs)
    eng = ll.Engine(cfg, sizes=sizes, population=pop, top=4, precision=32, mode=1, device=0)
    eng.set_members(0, members)
    t0 = time.time()
    fit, lr, steps, st = eng.evaluate(spad, lpad, d, 5000)
    dt = time.time() - t0
    ep_steps, flags, ep_fit = eng.read_episodes()
    res[name] = (fit.copy(), steps.copy(), ep_steps.copy(), flags.copy())
    print(name, eng.info.kernel.decode(), "steps", steps.sum(), "success rate", (flags & 4).astype(bool).mean(), "mean fit", fit.mean(), f"{dt:.3f}s", flush=True)
    eng.close()
want = po.evaluate(cfg.env_cfg(), sizes, 1, members.astype(np.float32), spad, lpad, d, 5000, 32, threads=8)
print("oracle fp32 exact: steps", want["steps"].sum(), "success", (want["ep_flags"] & 4).astype(bool).mean(), "mean fit", want["fitness"].mean())
w_ep = want["ep_steps"] if "ep_steps" in want else None
for name, (fit, steps, ep, fl) in res.items():
    line = f"{name}: member steps == oracle {np.mean(steps == want['steps']):.3f}"
    if w_ep is not None:
        line += f" | episode steps == oracle {np.mean(ep == w_ep):.3f}"
    line += f" | episode steps == ffma {np.mean(ep == res['ffma'][2]):.3f} | median rel member-step diff {np.median(np.abs(steps - want['steps']) / want['steps']):.5f}"
    print(line)
