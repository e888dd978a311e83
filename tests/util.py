"""Shared helpers of the test suite (fixtures loading, synthetic populations)."""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
CFG_PATH = os.path.join(ROOT, "config.txt")

TOPOLOGIES = {
    "16_16": [5, 16, 16, 4], "16_32": [5, 16, 32, 4], "8_16_8": [5, 8, 16, 8, 4], "16_64_32": [5, 16, 64, 32, 4],
    "32_128_32": [5, 32, 128, 32, 4], "8_16_32_16_8": [5, 8, 16, 32, 16, 8, 4],
}


def param_count(sizes):
    return sum(sizes[l] * sizes[l + 1] + sizes[l + 1] for l in range(len(sizes) - 1))


def golden():
    with open(os.path.join(GOLDEN, "ref_outputs.json")) as f:
        return json.load(f)


def elites(topo):
    z = np.load(os.path.join(GOLDEN, f"elites_{topo}.npz"))
    return z["params"], [int(s) for s in z["sizes"]]


def layouts(n):
    g = golden()["layouts"][str(n)]
    return (np.array(g["spad_x1"], dtype=np.int32), np.array(g["lpad_x1"], dtype=np.int32), np.array(g["directions"], dtype=np.uint64))


def constant_action_member(sizes, action):
    """All-zero network whose last-layer bias makes `action` the first maximum (canonical parameter order)."""
    P = param_count(sizes)
    p = np.zeros(P)
    last_b = P - sizes[-1]
    p[last_b + action] = 1.0
    return p


def random_population(sizes, pop, seed, scale=1.0):
    rng = np.random.default_rng(seed)
    return rng.normal(scale=scale, size=(pop, param_count(sizes)))


def elite_population(topo, pop, seed, noise=0.05):
    """S-elite: member j = elite[j mod 10] + N(0, noise^2)."""
    e, sizes = elites(topo)
    rng = np.random.default_rng(seed)
    p = e[np.arange(pop) % e.shape[0]] + rng.normal(scale=noise, size=(pop, e.shape[1]))
    return p, sizes


def cfg_for(topo_sizes, **overrides):
    import lunar_lander_b200 as ll
    ov = {"NNConfig.hlayers": ",".join(str(s) for s in topo_sizes[1:-1])}
    ov.update(overrides)
    return ll.Config(CFG_PATH, ov)


def injected_mask(pop, top, ratio):
    """Child slots that hold fresh random individuals (ll_is_injected, csrc/ll_philox.h): n_inject of the pop - top children,
    spread evenly over the child slots."""
    n_children = pop - top
    n_inject = min(n_children, int(np.floor(ratio * n_children + 0.5)))
    c = np.arange(n_children, dtype=np.int64)
    m = np.zeros(pop, dtype=bool)
    m[top:] = ((c + 1) * n_inject) // n_children != (c * n_inject) // n_children
    assert m.sum() == n_inject
    return m
