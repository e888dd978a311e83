"""Generates the committed fixtures under tests/golden/ — run HERE (the build container), where /root/reference exists:

    python tests/golden/make_golden.py

Sources of truth:
  * sample weight files of the reference (sample_runs/*/lunar_lander_*.txt) -> elites_<topology>.npz (top-10 members,
    canonical order: per layer W[out][in] row-major then b[out], float64 as parsed from the 6-digit text), plus one
    file copied verbatim as a format fixture;
  * the reference's OWN sources compiled unchanged (oracle/_ref/libll_ref.so; oracle/Makefile): constant-action
    episodes (no network involved -> reference-owned known answers), the seed-5247 layout set
    after balance_layouts, and per-layout outcomes of elite members driven by the restated policy network
    (reference environment + restated forward pass -> regression anchors, "parity unpinned" at the ma-libs boundary).
Nothing under tests/golden is read from /root/reference at test time.
"""
import json
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import pyoracle as po  # noqa: E402

REF = "/root/reference"
CFG = os.path.join(ROOT, "config.txt")

FILES = {
    "16_16": "sample_runs/16_16/lunar_lander_5_16_16_4_s_6500.txt",
    "8_16_8": "sample_runs/8_16_8/lunar_lander_5_8_16_8_4_s_0500_sample_2.txt",
    "16_64_32": "sample_runs/16_64_32/lunar_lander_5_16_64_32_4_s_1100.txt",
    "32_128_32": "sample_runs/32_128_32/lunar_lander_5_32_128_32_4_s_0800.txt",
    "8_16_32_16_8": "sample_runs/8_16_32_16_8/lunar_lander_5_8_16_32_16_8_4_s_2500.txt",
}
# values in which the sample configs differ from config.txt, per topology
CFG_EXTRA = {"8_16_8": {"NNTrainingCfg.tp_no_fuel_penalty": 500.0}}


def parse_weight_file(path):
    """Independent pure-Python reader of the App. B format (does not use the product's or the oracle's reader)."""
    tok = open(path).read().split()
    i = 0
    hdr = {}
    while tok[i] != "vSize":
        hdr[tok[i]] = tok[i + 1]
        i += 2
    n_layers = int(hdr["nLayers"])
    sizes = [int(t) for t in tok[i + 1:i + 1 + n_layers]]
    i += 1 + n_layers
    top = int(hdr["nTop"])
    P = sum(sizes[l] * sizes[l + 1] + sizes[l + 1] for l in range(n_layers - 1))
    w_at, b_at, at = [], [], 0
    for l in range(n_layers - 1):
        w_at.append(at)
        b_at.append(at + sizes[l] * sizes[l + 1])
        at = b_at[-1] + sizes[l + 1]
    params = np.zeros((top, P))
    for section, start, end in ((0, "vBiases_start", "vBiases_end"), (1, "vWeights_start", "vWeights_end")):
        assert tok[i] == start, (tok[i], start)
        i += 1
        while tok[i] != end:
            assert tok[i] == "pop" and tok[i + 2] == "layer" and tok[i + 4] == "dims"
            m, l, rows, cols = int(tok[i + 1]), int(tok[i + 3]), int(tok[i + 5]), int(tok[i + 6])
            i += 7
            vals = [float(t) for t in tok[i:i + rows * cols]]
            i += rows * cols
            base = b_at[l] if section == 0 else w_at[l]
            params[m, base:base + rows * cols] = vals
        i += 1
    assert tok[i] == "generator"
    return hdr, sizes, params, len(tok) - i - 1


def main():
    assert po.ref() is not None, "build oracle/_ref first (make -C oracle ref)"
    out = {"env_known_answers": [], "layouts": {}, "elite_runs": {}}

    # 1. constant-action episodes through the reference environment (reference-owned known answers)
    for prec in (64, 32):
        for action in (0, 1, 2, 3):
            r = po.ref_trace(CFG, [5, 16, 16, 4], 1, None, 402, 131, 0, 5000, prec, fixed_action=action, cap=5000)
            last = r["rec"][-1]
            out["env_known_answers"].append(dict(precision=prec, action=action, spad_x1=402, lpad_x1=131, max_steps=5000,
                                                 steps=r["steps"], flags=r["flags"], step_penalty_sum=r["acc"], terminal_penalty=r["term"],
                                                 final=[float(v) for v in last[6:11]], final_state=[float(v) for v in last[1:6]]))

    # 2. the layout set of train(): seeds 5247+i, balanced to 25/25 (src/nn_engine_trainer.cpp:283-309, :196-260)
    for n in (1, 8, 50):
        spad, lpad, d = po.ref_make_layouts(CFG, 5247, n)
        out["layouts"][str(n)] = dict(spad_x1=spad.tolist(), lpad_x1=lpad.tolist(), directions=d.tolist())

    # 3. elites of the sample runs and their behaviour under reference env + restated policy
    spad, lpad, d = po.ref_make_layouts(CFG, 5247, 50)
    for topo, rel in FILES.items():
        hdr, sizes, params, n_gen = parse_weight_file(os.path.join(REF, rel))
        np.savez_compressed(os.path.join(HERE, f"elites_{topo}.npz"), params=params, sizes=np.array(sizes),
                            epochs=int(hdr["nEpochs"]), top=int(hdr["nTop"]), pop=int(hdr["nPopSize"]), act=int(hdr["act"]),
                            ratio=float(hdr["randomInjectionRatio"]), generator_words=n_gen)
        extra = CFG_EXTRA.get(topo)
        runs = {}
        for prec in (64, 32):
            ev = po.ref_evaluate(CFG, sizes, int(hdr["act"]), params[:2], spad, lpad, d, 5000, prec, extra=extra)
            per_layout = []
            for m in range(2):
                rows = []
                for l in range(50):
                    t = po.ref_trace(CFG, sizes, int(hdr["act"]), params[m], spad[l], lpad[l], int(d[l]), 5000, prec, cap=0, extra=extra)
                    rows.append([t["steps"], t["flags"], t["term"] + t["acc"]])
                per_layout.append(rows)
            runs[str(prec)] = dict(fitness=ev["fitness"].tolist(), lr=ev["lr"].tolist(), steps=ev["steps"].tolist(), per_layout=per_layout)
        out["elite_runs"][topo] = dict(file=rel, extra=extra or {}, runs=runs)
        print(topo, sizes, "member0 f64:", runs["64"]["fitness"][0], runs["64"]["steps"][0],
              "landed", sum(1 for r in runs["64"]["per_layout"][0] if r[1] & 4))

    shutil.copyfile(os.path.join(REF, FILES["8_16_8"]), os.path.join(HERE, "sample_lunar_lander_5_8_16_8_4_s_0500.txt"))
    with open(os.path.join(HERE, "ref_outputs.json"), "w") as f:
        json.dump(out, f, indent=0)
    print("written", os.path.join(HERE, "ref_outputs.json"))


if __name__ == "__main__":
    main()
