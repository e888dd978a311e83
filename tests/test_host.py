"""CPU tests (-m "not gpu") of everything that runs without a device: the C-ABI library loads and exports every
symbol include/ll_gpu.h declares, fails loudly without a GPU, the host-side helpers (config, weight-file header,
candidate merge, sharding arithmetic) and the two-rank exchange logic over gloo."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import util


@pytest.fixture(scope="module")
def ll(built):
    import lunar_lander_b200 as m
    return m


def test_abi_exports_every_declared_symbol(ll):
    hdr = open(os.path.join(util.ROOT, "include", "ll_gpu.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(ll_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 40
    from lunar_lander_b200 import _abi
    lib = C.CDLL(_abi.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), f"libll_gpu.so does not export {name}"
    assert declared == set(_abi.SIGNATURES), declared ^ set(_abi.SIGNATURES)
    assert lib.ll_abi_version() == 1
    out = subprocess.run(["nm", "-D", "--defined-only", _abi.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    assert declared <= exported


def test_no_cpu_fallback(ll):
    """Without a CUDA device the product refuses to run (there is no CPU path behind the ABI)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    cfg = ll.Config(util.CFG_PATH)
    with pytest.raises(ll.LLError) as e:
        ll.Engine(cfg, population=8, top=2)
    assert e.value.code == 2 and "no CUDA device" in str(e.value)
    from lunar_lander_b200 import _abi
    v = C.c_double()
    assert _abi.lib.ll_microbench_fma(0, 32, 0.01, C.byref(v), None) == 2


def test_cpp_host_trainer_refuses_without_gpu(built, tmp_path):
    """main_train_d (the C++ mirror of src/main_train.cpp) reads config.txt like the reference and, with no CUDA device,
    exits 1 with the library's message instead of training on the CPU."""
    import shutil
    import torch
    exe = os.path.join(util.ROOT, "lunar_lander_b200", "host", "bin", "main_train_d")
    assert os.path.exists(exe)
    rc = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True)
    assert rc.returncode == 1 and "FATAL: Could not load configuration from config.txt" in rc.stderr
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    shutil.copy(util.CFG_PATH, tmp_path / "config.txt")
    rc = subprocess.run([exe], cwd=tmp_path, capture_output=True, text=True)
    assert rc.returncode == 1 and "Using data type: double" in rc.stdout
    assert "no CUDA device" in rc.stderr and "Failed to initialize NNEngineTrainer." in rc.stderr
    out = subprocess.run(["ldd", exe], capture_output=True, text=True).stdout
    assert "libll_gpu.so" in out and "oracle" not in out


def test_python_drop_in_modules_import_by_their_top_level_names(built, tmp_path):
    """with lunar_lander_b200/pymodules on sys.path (and nothing else of this repo) a fresh interpreter imports
    `lunar_lander_cpp` and `cpp_nn_py` the way main.py:24-28 and mod_nn_play.py:20 do; names and defaults are the reference's
    (src/lunar_lander_cpp.cpp:21-106, mod_nn_play.py:28-56); without a GPU construction fails loudly, nothing computes on the CPU."""
    import json
    import torch
    code = r'''
import json, sys
import numpy as np
import lunar_lander_cpp, cpp_nn_py
from lunar_lander_cpp import GameLogicCppDouble as GameLogicCpp
from lunar_lander_cpp import generate_random_pad_positions
out = {"files": [lunar_lander_cpp.__file__, cpp_nn_py.__file__]}
out["pads"] = list(generate_random_pad_positions(960.0, 80, 200, seed=5247))
out["pads_kw"] = list(generate_random_pad_positions(screen_width=960.0, spad_width=80, lpad_width=200, seed=None, force_left_to_right=True, force_right_to_left=False))
for cls in (lunar_lander_cpp.GameLogicCppDouble, lunar_lander_cpp.GameLogicCppFloat):
    for name in ("x", "y", "vx", "vy", "fuel", "landed", "crashed", "landed_successfully", "landing_pad_center_x", "landing_pad_y", "no_print",
                 "last_action", "reset", "is_done", "get_state", "update", "calculate_step_penalty", "calculate_terminal_penalty",
                 "calculate_terminal_penalty_lr_bonus"):
        assert hasattr(cls, name), name
for cls in (cpp_nn_py.ANN_MLP_GA_double, cpp_nn_py.ANN_MLP_GA_float):
    for name in ("SetName", "Deserialize", "Serialize", "GetEpochs", "GetNetworkSize", "GetPopSize", "GetTopPerformersSize", "feedforwardIndex"):
        assert hasattr(cls, name), name
try:
    GameLogicCpp(False, "no_such_config.txt")
    out["missing_cfg"] = "no error"
except RuntimeError as e:
    out["missing_cfg"] = str(e)
try:
    GameLogicCpp(True, sys.argv[1])
    out["ctor"] = "ok"
except RuntimeError as e:
    out["ctor"] = str(e)
net = cpp_nn_py.ANN_MLP_GA_double()
net.SetName("lunar_lander")
out["epochs_empty"] = net.GetEpochs()
try:
    net.feedforwardIndex(np.zeros(5), 0)
    out["ff_empty"] = "no error"
except RuntimeError as e:
    out["ff_empty"] = str(e)
print(json.dumps(out))
'''
    pymods = os.path.join(util.ROOT, "lunar_lander_b200", "pymodules")
    env = {k: v for k, v in os.environ.items() if k != "PYTHONPATH"}
    env["PYTHONPATH"] = pymods
    p = subprocess.run([sys.executable, "-c", code, util.CFG_PATH], cwd=tmp_path, capture_output=True, text=True, env=env, timeout=300)
    assert p.returncode == 0, p.stderr
    out = json.loads(p.stdout.strip().splitlines()[-1])
    assert all(os.path.dirname(f) == pymods and f.endswith(".so") for f in out["files"])
    assert out["pads"] == [402, 131]  # libstdc++ mt19937(5247) + uniform_int_distribution, SURVEY.md App. D
    assert out["pads_kw"][0] + 40 < out["pads_kw"][1] + 100
    assert out["missing_cfg"] == "FATAL: Could not load configuration from no_such_config.txt. Exiting."  # src/game_logic.cpp:44
    assert out["epochs_empty"] == 0 and "no shape" in out["ff_empty"]
    if torch.cuda.is_available():
        assert out["ctor"] == "ok"
    else:
        assert "no CUDA device" in out["ctor"]


def test_product_does_not_link_the_oracle(ll):
    from lunar_lander_b200 import _abi
    out = subprocess.run(["ldd", _abi.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "ll_ref" not in out
    for root, _, files in os.walk(os.path.join(util.ROOT, "lunar_lander_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".h", ".cuh")):
                src = open(os.path.join(root, f)).read()
                assert "pyoracle" not in src and "libll_oracle" not in src and "oracle/_ref" not in src, f


def test_config_reader(ll, tmp_path):
    cfg = ll.Config(util.CFG_PATH)
    assert cfg.getInt("NNConfig.population_size") == 100 and cfg.getInt("NNConfig.top_individuals") == 10
    assert cfg.getVectorInt("NNConfig.hlayers") == [16, 16] and cfg.layer_sizes() == [5, 16, 16, 4]
    assert cfg.getBool("NNConfig.elitism") is True and cfg.getString("NNConfig.name") == "lunar_lander"
    assert cfg.getVectorDouble("GameCfg.x0") == [52.0, 433.8] and cfg.getDouble("NNConfig.random_injection_ratio") == 0.35
    env = cfg.env_cfg()
    assert env.max_fuel == 1000.0 and env.tp_landed_lr_bonus == 10000.0 and env.y0 == 433.8
    p = tmp_path / "c.txt"
    p.write_text("A.b = 3  # comment\nUnknown.key = whatever\nA.v = 1, 2,3\n# only a comment\n")
    c2 = ll.Config(str(p))
    assert c2.getInt("A.b") == 3 and c2.getVectorInt("A.v") == [1, 2, 3]
    with pytest.raises(RuntimeError):
        ll.Config(str(tmp_path / "missing.txt"))


def test_weight_file_header(ll):
    hdr = ll.peek_weight_file(os.path.join(util.GOLDEN, "sample_lunar_lander_5_8_16_8_4_s_0500.txt"))
    assert hdr == dict(sizes=[5, 8, 16, 8, 4], population=100, top=10, epochs=500, act=1, strategy=2, ratio=hdr["ratio"], name="lunar_lander")
    with pytest.raises(ll.LLError) as e:
        ll.peek_weight_file("/nonexistent/file.txt")
    assert e.value.code == 4


def test_layout_helpers_error_paths(ll):
    from lunar_lander_b200 import _abi
    s, l = C.c_int32(), C.c_int32()
    assert _abi.lib.ll_generate_random_pad_positions(100.0, 80, 200, 1, 1, 0, 0, C.byref(s), C.byref(l)) == 1  # can never fit
    a, b = ll.generate_random_pad_positions(960, 80, 200, 7, force_left_to_right=True)
    assert a + 40 < b + 100
    a, b = ll.generate_random_pad_positions(960, 80, 200, 7, force_right_to_left=True)
    assert a + 40 > b + 100
    # balance_layouts' early return leaves the directions untouched
    cfg = ll.Config(util.CFG_PATH)
    stale = np.full(2, 1, dtype=np.uint64)
    spad, lpad, d, sc, lc = ll.make_layouts(cfg, 5247, 2, directions=stale)
    n_ltr = int((sc < lc).sum())
    if n_ltr == 1:
        assert d.tolist() == [1, 1]


def test_merge_candidates_matches_global_order():
    from lunar_lander_b200.sharded import merge_candidates, pad_candidates
    rng = np.random.default_rng(3)
    fit = np.round(rng.normal(size=1000), 1)  # many ties
    fit[17] = -0.0
    fit[5] = 0.0
    top, world = 16, 4
    blocks = []
    for r in range(world):
        lo, hi = 1000 * r // world, 1000 * (r + 1) // world
        f = fit[lo:hi]
        order = np.lexsort((np.arange(lo, hi), f + 0.0))[:top]
        blocks.append(pad_candidates(f[order], (order + lo).astype(np.uint64), top))
    f, idx = merge_candidates(np.concatenate(blocks), top)
    want = np.lexsort((np.arange(1000), fit + 0.0))[:top]
    assert idx.tolist() == want.tolist() and np.array_equal(f, fit[want] + 0.0)
    short = pad_candidates(np.array([1.0]), np.array([3], dtype=np.uint64), 4)
    assert short[1:, 0].tolist() == [np.inf] * 3


def _gloo_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from lunar_lander_b200.sharded import merge_candidates, pad_candidates
    rng = np.random.default_rng(11)
    pop, top, P = 501, 8, 12
    fit = np.round(rng.normal(size=pop), 1)
    params = rng.normal(size=(pop, P))
    lo, hi = pop * rank // world, pop * (rank + 1) // world
    order = np.lexsort((np.arange(lo, hi), fit[lo:hi] + 0.0))[:top]
    local = torch.from_numpy(pad_candidates(fit[lo:hi][order], (order + lo).astype(np.uint64), top))
    gathered = torch.empty((world * top, 2), dtype=torch.float64)
    dist.all_gather_into_tensor(gathered, local)
    _, elite = merge_candidates(gathered.numpy(), top)
    # elite rows: owners contribute bits, others zeros; integer SUM == broadcast
    buf = np.zeros((top, P))
    for r, g in enumerate(elite):
        if lo <= g < hi:
            buf[r] = params[int(g)]
    t = torch.from_numpy(buf.view(np.int64).copy())
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    got = t.numpy().view(np.float64)
    want = np.lexsort((np.arange(pop), fit + 0.0))[:top]
    ok = elite.tolist() == want.tolist() and np.array_equal(got, params[want])
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_two_rank_exchange_over_gloo():
    """world_size-2 run of the per-generation exchange (candidate all-gather + elite all-reduce) on CPU."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]


def test_philox_known_answers(built):
    """Philox4x32-10 known-answer vectors of the Random123 distribution (Salmon et al., SC'11)."""
    import pyoracle as po
    out = (C.c_uint32 * 4)()
    kat = [((0, 0, 0, 0), (0, 0), "6627e8d5 e169c58d bc57ac4c 9b00dbd8"),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, "408f276d 41c83b0e a20bc7c6 6d5451fd"),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), "d16cfe09 94fdcceb 5001e420 24126ea1")]
    for c, k, want in kat:
        po.port().orc_philox(*[C.c_uint32(x) for x in c], *[C.c_uint32(x) for x in k], out)
        assert " ".join(f"{x:08x}" for x in out) == want


def test_create_population_statistics(built):
    import pyoracle as po
    sizes = util.TOPOLOGIES["8_16_8"]
    P = util.param_count(sizes)
    pop, top = 400, 10
    p0 = po.create_population(sizes, 5247, pop, top, 0, 2, 0.35, True, False, 1.0, 0.1, np.zeros((pop, P)))
    assert abs(p0.mean()) < 0.01 and abs(p0.std() - 1.0) < 0.01
    p1 = po.create_population(sizes, 5247, pop, top, 1, 2, 0.35, True, True, 1.0, 0.1, p0)
    assert np.array_equal(p1[:top], p0[:top])
    fresh = util.injected_mask(pop, top, 0.35)  # spread evenly over the child slots: any contiguous shard gets the same mix
    assert fresh[top:top + 100].sum() in (34, 35, 36)
    kids = p1[top:][~fresh[top:]]
    d = np.abs(kids[:, None, :] - p0[None, :top, :]).mean(axis=2).min(axis=1)  # distance to the nearest elite
    assert d.max() < 0.15 and abs(p1[fresh].std() - 1.0) < 0.02
    # generation / seed / member all change the draw; the same key reproduces it
    assert not np.array_equal(p1, po.create_population(sizes, 5247, pop, top, 2, 2, 0.35, True, True, 1.0, 0.1, p0))
    assert np.array_equal(p1, po.create_population(sizes, 5247, pop, top, 1, 2, 0.35, True, True, 1.0, 0.1, p0))


def test_float_population_draws_are_standard_normal(built):
    """fp32 populations draw their Gaussians in float (ll_normal4_f32: Philox block -> two Box-Muller pairs with deterministic float
    log / sin / cos, csrc/ll_philox.h): moments, tails and independence of the four variates of a block."""
    import pyoracle as po
    from scipy import stats
    sizes = util.TOPOLOGIES["16_64_32"]
    P = util.param_count(sizes)
    pop = 300
    p0 = po.create_population(sizes, 5247, pop, 10, 0, 2, 0.35, True, False, 1.0, 0.1, np.zeros((pop, P)), 32).astype(np.float64)
    x = p0.ravel()
    assert abs(x.mean()) < 4.0 / np.sqrt(x.size) and abs(x.std() - 1.0) < 5e-3
    assert abs(stats.skew(x)) < 0.02 and abs(stats.kurtosis(x)) < 0.03
    assert stats.kstest(x[:200000], "norm").pvalue > 1e-3
    assert np.abs(x).max() < 6.5 and (np.abs(x) > 4).mean() < 2e-4          # radius from 23-bit uniforms: |x| <= sqrt(-2 ln 2^-24) = 5.77
    c = np.corrcoef(p0[:, :-1].ravel(), p0[:, 1:].ravel())[0, 1]          # neighbouring parameters (same Philox block or the next)
    assert abs(c) < 0.01
    p1 = po.create_population(sizes, 5247, pop, 10, 3, 2, 0.35, True, True, 1.0, 0.1, p0, 32)
    fresh = util.injected_mask(pop, 10, 0.35)
    kids = (p1[10:][~fresh[10:]]).astype(np.float64)
    d = np.abs(kids[:, None, :] - p0[None, :10, :]).mean(axis=2).min(axis=1)
    assert d.max() < 0.15 and d.min() > 0.05                               # children = parent + N(0, 0.1^2): E|d| = 0.08


def test_bench_reference_arm_prints_the_contract_line(built):
    """`bench.py --impl reference` (the driver's reference arm: the reference's own train_generation, oracle/_ref, on the host cores)
    prints one JSON line with the contract's keys — runs without a GPU."""
    import json
    import subprocess
    import sys
    p = subprocess.run([sys.executable, os.path.join(util.ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1", "--ref-step-seconds", "1"],
                       capture_output=True, text=True, timeout=600, cwd=util.ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    line = json.loads(p.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "lander-steps/sec" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["unit"] == "lander-steps/s" and line["steps"] == 1 and line["n_gpus"] == 1
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1 and line["cpu_baseline"]["sample"]
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert line["config"]["workload"].startswith("configs[3]")


def test_bench_flop_accounting():
    """The per-step flop figures the bench line's roofline and tensor_pipe keys are built from."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(util.ROOT, "bench.py"))
    b = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(b)
    assert b.tc05_flop_per_step([5, 32, 128, 32, 4]) == 52736  # (2 x 32 + 6 x 128 + 24 x 32 + 6 x 8) x 16 x 2
    sizes = [5, 32, 128, 32, 4]
    assert sum(2 * sizes[i] * sizes[i + 1] + sizes[i + 1] for i in range(4)) == 17156  # 2 MAC + N, BASELINE.md section 3
