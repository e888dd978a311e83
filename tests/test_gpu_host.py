"""The C++ host mirror (lunar_lander_b200/host: Config, nn::ANN_MLP_GA<T>, Training::NNEngineTrainer<T>, main_train_d/_f)
against the reference's own main_train compiled unchanged (oracle/_ref/main_train_d/_f, built from /root/reference/src by
oracle/Makefile with the ma-libs stand-ins): same config.txt, same seeds ->
  * identical console log (generation blocks, best member, fitness to 4 decimals, stagnation resets) except wall times,
  * identical checkpoint files up to the `generator` line, identical file names,
  * identical behaviour of --continue (resume from <name>_last.txt) and of the unreachable --step file name.
Reference interfaces: src/main_train.cpp:14-106, src/nn_engine_trainer.cpp:67-438.

What this identity pins, and what it does not: both binaries link a stand-in for the un-vendored ma-libs (oracle/shim: config loader,
thread pool, nn::ANN_MLP_GA<T>), and that stand-in includes the product's ll_math.h / ll_philox.h.  So the identity holds MODULO THE
MA-LIBS SHIM: it pins the reference's host loop, environment, penalties, layout generation / balancing, logging and file naming
(all compiled from the reference's own sources); the policy forward pass, the GA and the Serialize token format are this repo's
restatement on both sides (parity unpinned at the ma-libs boundary; tests/test_oracle.py measures how the
restated tanh flies the sample elites against libm's).
"""
import os
import re
import subprocess

import pytest

from util import CFG_PATH, ROOT

BIN = os.path.join(ROOT, "lunar_lander_b200", "host", "bin")
REF = os.path.join(ROOT, "oracle", "_ref")

pytestmark = pytest.mark.gpu


def write_config(dirname, **ov):
    lines = []
    with open(CFG_PATH) as f:
        for line in f:
            key = line.split("=", 1)[0].strip() if "=" in line else None
            if key in ov:
                line = f"{key} = {ov.pop(key)}\n"
            lines.append(line)
    for k, v in ov.items():
        lines.append(f"{k} = {v}\n")
    os.makedirs(dirname, exist_ok=True)
    with open(os.path.join(dirname, "config.txt"), "w") as f:
        f.writelines(lines)


def run(binary, cwd, *args):
    p = subprocess.run([binary, *args], cwd=cwd, capture_output=True, text=True, timeout=600)
    return p.returncode, p.stdout, p.stderr


def normalise(log):
    """drop the only run-dependent part of the log: wall times"""
    return re.sub(r"\(Took: \d+:\d+, Total: \d+:\d+\)", "(Took: T, Total: T)", log)


def weight_file_body(path):
    with open(path) as f:
        text = f.read()
    head, sep, _ = text.partition("generator")
    assert sep, f"{path}: no generator line"
    return head


def both(tmp_path, suffix, args=(), gpu_args=(), **ov):
    ours, ref = str(tmp_path / "ours"), str(tmp_path / "ref")
    for d in (ours, ref):
        if not os.path.exists(os.path.join(d, "config.txt")):
            write_config(d, **ov)
    ref_bin = os.path.join(REF, "main_train_" + suffix)
    if not os.path.exists(ref_bin):
        pytest.skip("oracle/_ref was not built (needs /root/reference at build time)")
    r_ref = run(ref_bin, ref, *args)
    r_our = run(os.path.join(BIN, "main_train_" + suffix), ours, *args, *gpu_args)
    return ours, ref, r_our, r_ref


def compare_runs(ours, ref, r_our, r_ref):
    assert r_ref[0] == 0, r_ref[2]
    assert r_our[0] == 0, r_our[2]
    assert normalise(r_our[1]) == normalise(r_ref[1])
    files_ref = sorted(os.listdir(os.path.join(ref, "data")))
    files_our = sorted(os.listdir(os.path.join(ours, "data")))
    assert files_our == files_ref and files_ref
    for name in files_ref:
        assert weight_file_body(os.path.join(ours, "data", name)) == weight_file_body(os.path.join(ref, "data", name)), name
    return files_ref


SMALL = {"NNConfig.epochs": 12, "NNConfig.population_size": 48, "NNTrainingCfg.layout_nb": 6, "NNConfig.save_interval": 5,
         "NNTrainingCfg.reset_period": 2}


def test_main_train_d_matches_reference_log_and_files(tmp_path, built):
    ours, ref, r_our, r_ref = both(tmp_path, "d", **SMALL)
    files = compare_runs(ours, ref, r_our, r_ref)
    assert "lunar_lander_last.txt" in files and "lunar_lander_5_16_16_4_s_0010.txt" in files
    assert "Stagnation detected" in r_ref[1], "the case is meant to cover a layout reset"
    assert "ll_episode_exactw_kernel<double>" in r_our[2]  # 100 x 50: the small-population exact kernel


def test_continue_resumes_like_the_reference(tmp_path, built):
    ours, ref, r_our, r_ref = both(tmp_path, "d", **SMALL)
    compare_runs(ours, ref, r_our, r_ref)
    ours, ref, r_our, r_ref = both(tmp_path, "d", args=("--continue",))
    files = compare_runs(ours, ref, r_our, r_ref)
    assert "Successfully loaded network from checkpoint 'last'." in r_our[1]
    assert "--- Generation 13 ---" in r_our[1] and "lunar_lander_5_16_16_4_s_0024.txt" in files


def test_deep_topology_host_select(tmp_path, built):
    """BASELINE configs[4]'s topology (8_16_32_16_8) through the trainer, host std::sort selection: identical to the reference"""
    ov = dict(SMALL)
    ov.update({"NNConfig.hlayers": "8,16,32,16,8", "NNConfig.epochs": 6, "NNConfig.population_size": 96})
    ours, ref, r_our, r_ref = both(tmp_path, "d", gpu_args=("--set", "GpuCfg.select=host"), **ov)
    compare_runs(ours, ref, r_our, r_ref)


def test_device_select_differs_only_in_tie_order(tmp_path, built):
    """GpuCfg.select = device ranks by (fitness, index); the reference's unstable std::sort may order members of EQUAL fitness
    differently (random nets that repeat one action fly identical trajectories), which can swap a low-ranked parent.  The
    best member, its fitness and the landing count per generation are the reference's; the population mean stays close."""
    ov = dict(SMALL)
    ov.update({"NNConfig.hlayers": "8,16,32,16,8", "NNConfig.epochs": 6, "NNConfig.population_size": 96})
    ours, ref, r_our, r_ref = both(tmp_path, "d", gpu_args=("--set", "GpuCfg.select=device"), **ov)
    assert r_our[0] == 0 and r_ref[0] == 0, r_our[2]
    pick = lambda log, pat: re.findall(pat, log)
    for pat in (r"Best Avg Fitness: (\S+) \(Member (\d+)\)", r"Both Directions: (\d+)", r"--- (Generation \d+|Stagnation[^-]*) ---"):
        assert pick(r_our[1], pat) == pick(r_ref[1], pat) and pick(r_ref[1], pat)
    mean_our = [float(x) for x in re.findall(r"^  Avg Fitness:\s+(\S+)", r_our[1], re.M)]
    mean_ref = [float(x) for x in re.findall(r"^  Avg Fitness:\s+(\S+)", r_ref[1], re.M)]
    assert len(mean_our) == 6 and all(abs(a - b) <= 0.05 * abs(b) for a, b in zip(mean_our, mean_ref))


def test_main_train_f_exact_mode_matches_reference(tmp_path, built):
    ours, ref, r_our, r_ref = both(tmp_path, "f", gpu_args=("--set", "GpuCfg.mode=exact"), **SMALL)
    compare_runs(ours, ref, r_our, r_ref)
    assert "Using data type: float" in r_our[1]


def test_main_train_f_fast_mode_trains(tmp_path, built):
    """default float build = throughput mode (tensor-core kernel): not bit-comparable, must run and improve or hold the best"""
    d = str(tmp_path / "fast")
    ov = dict(SMALL)
    ov.update({"NNConfig.hlayers": "16,64,32", "NNConfig.population_size": 512, "NNConfig.epochs": 8})
    write_config(d, **ov)
    rc, out, err = run(os.path.join(BIN, "main_train_f"), d)
    assert rc == 0, err
    best = [float(x) for x in re.findall(r"Best Avg Fitness: (-?[\d.]+)", out)]
    assert len(best) == 8
    assert "ll_episode_mmat_kernel" in err  # 6 layouts: the slots-on-N tensor-core kernel
    assert os.path.exists(os.path.join(d, "data", "lunar_lander_last.txt"))


def test_step_checkpoint_name_quirk_and_bad_arguments(tmp_path, built):
    """--step N looks for <name>_gen_000N.txt, a name save() never writes (src/main_train.cpp:48 vs nn_engine_trainer.cpp:161)"""
    d = str(tmp_path / "q")
    write_config(d, **dict(SMALL, **{"NNConfig.epochs": 1}))
    rc, out, err = run(os.path.join(BIN, "main_train_d"), d, "--step", "3")
    assert "Attempting to load network from checkpoint: gen_0003" in out
    assert "File not found" in err and "lunar_lander_gen_0003.txt" in err
    rc, out, err = run(os.path.join(BIN, "main_train_d"), d, "--step", "abc")
    assert rc == 1 and "not a valid integer" in err
    rc, out, err = run(os.path.join(BIN, "main_train_d"), d, "--step")
    assert rc == 1 and "requires an integer argument" in err
    rc, out, err = run(os.path.join(BIN, "main_train_d"), str(tmp_path))  # no config.txt there
    assert rc == 1 and "FATAL: Could not load configuration" in err


def test_python_sharded_trainer_prints_the_cpp_trainer_log(tmp_path, built):
    """lunar_lander_b200.trainer (the multi-GPU trainer: one process per GPU over torch.distributed, here with one process) against
    host/bin/main_train_d with GpuCfg.select = device: same generations, same best members and fitness, same files."""
    import sys
    ov = dict(SMALL)
    ov.update({"NNConfig.hlayers": "8,16,8", "NNConfig.epochs": 7, "NNConfig.population_size": 80})
    a, b = str(tmp_path / "cpp"), str(tmp_path / "py")
    write_config(a, **ov)
    write_config(b, **ov)
    r_cpp = run(os.path.join(BIN, "main_train_d"), a, "--set", "GpuCfg.select=device")
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    p = subprocess.run([sys.executable, "-m", "lunar_lander_b200.trainer"], cwd=b, capture_output=True, text=True, timeout=600, env=env)
    assert r_cpp[0] == 0 and p.returncode == 0, (r_cpp[2], p.stderr)
    assert normalise(p.stdout) == normalise(r_cpp[1])
    files = sorted(os.listdir(os.path.join(a, "data")))
    assert files == sorted(os.listdir(os.path.join(b, "data"))) and "lunar_lander_last.txt" in files
    for name in files:
        assert open(os.path.join(a, "data", name)).read() == open(os.path.join(b, "data", name)).read(), name


def _gpu_count():
    import torch
    return torch.cuda.device_count()


@pytest.mark.parametrize("suffix,mode,select", [("d", "exact", "auto"), ("f", "exact", "auto"), ("f", "fast", "auto"), ("f", "exact", "device")])
def test_cpp_trainer_on_two_gpus_prints_the_one_gpu_log(tmp_path, suffix, mode, select):
    """GpuCfg.gpus = 2: the C++ mirror shards the population over two GPUs (one context + one host thread per GPU, NCCL owned by the
    contexts: ll_comm_init, ll_select / ll_step_generation on shards).  Same config, same seeds -> the log and the checkpoint files
    of the one-GPU run, bit for bit (exact mode) — selection order, children (Philox keyed by global member), statistics.
    Runs only where two GPUs are visible (gpurun --gpus 2 and the driver's multi-GPU tier)."""
    if _gpu_count() < 2:
        pytest.skip("needs two GPUs")
    ov = {"NNConfig.hlayers": "16,64,32", "NNConfig.epochs": 6, "NNConfig.population_size": 999, "NNTrainingCfg.layout_nb": 8,
          "NNConfig.save_interval": 3, "GpuCfg.mode": mode, "GpuCfg.select": select}
    one, two = str(tmp_path / "one"), str(tmp_path / "two")
    write_config(one, **dict(ov, **{"GpuCfg.gpus": 1}))
    write_config(two, **dict(ov, **{"GpuCfg.gpus": 2}))
    r1 = run(os.path.join(BIN, "main_train_" + suffix), one)
    r2 = run(os.path.join(BIN, "main_train_" + suffix), two)
    assert r1[0] == 0, r1[2]
    assert r2[0] == 0, r2[2]
    banner = re.compile(r"^NCCL version .*\n", re.M)  # NCCL's own start-up line (printed when NCCL_DEBUG is set on the box)
    assert normalise(r1[1]) == banner.sub("", normalise(r2[1]))
    files = sorted(os.listdir(os.path.join(one, "data")))
    assert files == sorted(os.listdir(os.path.join(two, "data"))) and files
    for name in files:
        assert weight_file_body(os.path.join(one, "data", name)) == weight_file_body(os.path.join(two, "data", name)), name


def test_python_trainer_on_two_ranks_prints_the_one_rank_log(tmp_path, built):
    """torchrun --nproc-per-node 2 -m lunar_lander_b200.trainer (one process per GPU, the communicator behind the ABI: sharded.py
    broadcasts the NCCL id, ll_comm_init, ll_step_generation on shards) against the same trainer on one rank: same log, same
    checkpoint files, bit for bit (exact mode).  Runs only where two GPUs are visible."""
    import sys
    if _gpu_count() < 2:
        pytest.skip("needs two GPUs")
    ov = dict(SMALL)
    ov.update({"NNConfig.hlayers": "16,64,32", "NNConfig.epochs": 5, "NNConfig.population_size": 999, "NNTrainingCfg.layout_nb": 8,
               "NNConfig.save_interval": 2, "GpuCfg.mode": "exact"})
    one, two = str(tmp_path / "one"), str(tmp_path / "two")
    write_config(one, **ov)
    write_config(two, **ov)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    p1 = subprocess.run([sys.executable, "-m", "lunar_lander_b200.trainer", "--float"], cwd=one, capture_output=True, text=True, timeout=600, env=env)
    p2 = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                         "--master-port", "29533", "-m", "lunar_lander_b200.trainer", "--float"], cwd=two, capture_output=True, text=True, timeout=900, env=env)
    assert p1.returncode == 0, p1.stderr[-2000:]
    assert p2.returncode == 0, p2.stderr[-2000:]
    banner = re.compile(r"^NCCL version .*\n", re.M)
    assert normalise(p1.stdout) == banner.sub("", normalise(p2.stdout))
    files = sorted(os.listdir(os.path.join(one, "data")))
    assert files == sorted(os.listdir(os.path.join(two, "data"))) and files
    for name in files:
        assert weight_file_body(os.path.join(one, "data", name)) == weight_file_body(os.path.join(two, "data", name)), name
