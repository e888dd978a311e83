"""nn_play side of the boundary: the device-resident single lander (ll_sim_*, the GameLogicCpp<T>
object of main.py) behind the ``lunar_lander_cpp`` mirror, the ``cpp_nn_py`` mirror (mod_nn_play.py) and the headless
``main.py --mode=nn_play`` loop, each against the reference's own GameLogicCpp compiled unchanged (oracle/_ref) or the
port (oracle/ll_oracle.cpp).  Bit-exact: integer/flag results and every double/float state value.
"""
import os

import numpy as np
import pytest

import util
from util import CFG_PATH, ROOT

pytestmark = pytest.mark.gpu

SAMPLE = os.path.join(util.GOLDEN, "sample_lunar_lander_5_8_16_8_4_s_0500.txt")


@pytest.fixture(scope="module")
def mods(built):
    import pyoracle as po
    from lunar_lander_b200 import nn_play
    return po, nn_play.lunar_lander_cpp, nn_play.cpp_nn_py, nn_play


def oracle_trace(po, sizes, params, spad, lpad, direction, max_steps, precision, fixed_action=-1):
    """the compiled reference when it is there (this container / a box that received oracle/_ref), else the port"""
    if po.ref() is not None:
        return po.ref_trace(CFG_PATH, sizes, 1, params, spad, lpad, direction, max_steps, precision, fixed_action=fixed_action)
    import lunar_lander_b200 as ll
    return po.trace(ll.Config(CFG_PATH).env_cfg(), sizes, 1, params, spad, lpad, direction, max_steps, precision, fixed_action=fixed_action)


@pytest.mark.parametrize("precision", [64, 32])
@pytest.mark.parametrize("policy", ["elite", "noop", "up"])
def test_stepped_lander_equals_reference_game_logic(mods, precision, policy):
    """drive GameLogicCppDouble/Float with the action sequence of a reference episode: state after every update, done flag,
    public members, per-step penalty and both terminal-penalty overloads are the reference's, bit for bit"""
    po, llc, _, _ = mods
    params, sizes = util.elites("8_16_8")
    spad, lpad, dirs = util.layouts(8)
    fixed = {"elite": -1, "noop": 0, "up": 1}[policy]
    cls = llc.GameLogicCppDouble if precision == 64 else llc.GameLogicCppFloat
    dt = np.float64 if precision == 64 else np.float32
    logic = cls(True, CFG_PATH)
    for li in (0, 3):
        res = oracle_trace(po, sizes, params[0], int(spad[li]), int(lpad[li]), int(dirs[li]), 5000, precision, fixed)
        rec = res["rec"]
        logic.reset(float(spad[li]), float(lpad[li]))
        buf = np.zeros(5, dtype=dt)
        sp_sum, done = 0.0, False
        assert not logic.is_done() and logic.fuel == 1000.0 and logic.time_step == 0
        for t in range(res["steps"]):
            a = int(rec[t, 0])
            assert not done
            done = logic.update(a, buf)
            sp_sum += logic.calculate_step_penalty(a)
            assert np.array_equal(buf.astype(np.float64), rec[t, 1:6]), (li, t)
            assert (logic.x, logic.y, logic.vx, logic.vy, logic.fuel) == tuple(rec[t, 6:11]), (li, t)
            assert logic.last_action == a
        assert logic.time_step == res["steps"]
        flags = (1 if logic.landed else 0) | (2 if logic.crashed else 0) | (4 if logic.landed_successfully else 0)
        assert flags == res["flags"] and done == bool(res["flags"] & 3)
        assert sp_sum == res["acc"]
        term, bonus = logic.calculate_terminal_penalty_lr_bonus(res["steps"], int(dirs[li]))
        assert term == res["term"] and bonus == list(res["bonus"])
        assert logic.calculate_terminal_penalty(res["steps"]) == term  # the 1-argument overload computes the same value
        if done:  # update() on a finished game only re-emits the state (src/game_logic.cpp:297-301)
            before = (logic.x, logic.y, logic.fuel, logic.time_step)
            assert logic.update(1, buf) is True and before == (logic.x, logic.y, logic.fuel, logic.time_step)


def test_known_answers_constant_actions(mods):
    """noop crashes at step 9 (terminal 117.125); hold-UP times out at 5000 steps with an empty tank"""
    _, llc, _, _ = mods
    g = util.golden()
    logic = llc.GameLogicCppDouble(True, CFG_PATH)
    buf = np.zeros(5)
    ka = g["env_known_answers"][0]
    logic.reset(ka["spad_x1"], ka["lpad_x1"])
    n = 0
    while not logic.update(0, buf):
        n += 1
    assert n + 1 == 9 and logic.crashed and logic.landed and not logic.landed_successfully
    assert logic.calculate_terminal_penalty(9) == pytest.approx(117.125, abs=1e-12)
    logic.reset(ka["spad_x1"], ka["lpad_x1"])
    for _ in range(5000):
        done = logic.update(1, buf)
    assert not done and logic.fuel == 0.0
    assert logic.calculate_terminal_penalty(5000) == pytest.approx(3362.200000000099, rel=1e-15)
    assert g is not None


def test_binding_errors_and_members(mods, tmp_path):
    _, llc, _, _ = mods
    with pytest.raises(RuntimeError):
        llc.GameLogicCppDouble(True, str(tmp_path / "missing.txt"))  # src/game_logic.cpp:42-45
    logic = llc.GameLogicCppDouble(True, CFG_PATH)
    with pytest.raises(RuntimeError, match="1-dimensional"):
        logic.get_state(np.zeros((5, 1)))  # src/lunar_lander_cpp.cpp:50
    with pytest.raises(RuntimeError, match="1-dimensional"):
        logic.update(0, np.zeros((1, 5)))  # :63
    wrong = np.full(5, 7.0, dtype=np.float32)  # py::array_t<double> converts a float32 array into a temporary: the call succeeds
    logic.update(0, wrong)                     # and the caller's buffer stays untouched, exactly like the reference binding
    assert np.all(wrong == 7.0)
    logic.reset()
    with pytest.raises(AssertionError):
        logic.calculate_terminal_penalty_lr_bonus(10, 2)
    logic.reset(100, 500)
    assert logic.landing_pad_center_x == 600.0 and logic.landing_pad_y == 490.0
    assert logic.x == 100 + 40 - 65.5 / 2
    logic.fuel = 3.0  # def_readwrite
    logic.vx = 0.25
    buf = np.zeros(5)
    logic.get_state(buf)
    assert buf[4] == 3.0 / 1000 and buf[0] == 0.5 and logic.fuel == 3.0
    logic.reset()  # reset() keeps the pads of the last reset(spad, lpad) (src/game_logic.cpp:124-154)
    assert logic.landing_pad_center_x == 600.0 and logic.fuel == 1000.0
    s, l = llc.generate_random_pad_positions(960.0, 80, 200, seed=5247)
    g = util.golden()["layouts"]
    assert g is not None and 10 <= s <= 870 and 10 <= l <= 750
    s2, l2 = llc.generate_random_pad_positions(960.0, 80, 200, seed=5247, force_right_to_left=True)
    assert s2 + 40 > l2 + 100


def test_cpp_nn_py_mirror(mods):
    po, _, nnpy, _ = mods
    net = nnpy.ANN_MLP_GA_double()
    with pytest.raises(RuntimeError):
        net.feedforwardIndex(np.zeros(5), 0)
    with pytest.raises(RuntimeError):
        net.Deserialize("/nonexistent/file.txt")
    net.SetName("lunar_lander")
    net.Deserialize(SAMPLE)
    assert net.GetEpochs() == 500 and net.GetNetworkSize() == [5, 8, 16, 8, 4]
    assert net.GetPopSize() == 100 and net.GetTopPerformersSize() == 10
    params, sizes = util.elites("8_16_8")  # the same file's elites (tests/golden/make_golden.py)
    rng = np.random.default_rng(3)
    for _ in range(50):
        s = rng.uniform(-2, 2, 5)
        m = int(rng.integers(0, 10))
        assert net.feedforwardIndex(s, m) == po.forward(sizes, 1, params[m], s, 64)


@pytest.mark.parametrize("kw", [dict(seed=5), dict(seed=11, force_left_to_right=True), dict(seed=23, force_right_to_left=True),
                                dict(seed=1234)])
def test_headless_nn_play_equals_reference_loop(mods, kw):
    """main.py --mode=nn_play without the window: our GPU-backed loop vs the same loop over the reference's GameLogicCpp
    (oracle/ref_harness.cpp ref_nn_play): pads, start frame, every action and state, steps, outcome, penalty"""
    po, _, _, play = mods
    if po.ref() is None:
        pytest.skip("oracle/_ref was not built")
    r = play.nn_play(CFG_PATH, SAMPLE, record=True, log=lambda s: None, **kw)
    want = po.ref_nn_play(CFG_PATH, SAMPLE, seed=kw["seed"], force_ltr=kw.get("force_left_to_right", False),
                          force_rtl=kw.get("force_right_to_left", False), fps=60.0, max_frames=5000 + 180)
    assert (r.spad_x1, r.lpad_x1, r.frames, r.steps, r.forced_start_frame) == (
        want["spad_x1"], want["lpad_x1"], want["frames"], want["steps"], want["start_frame"])
    assert r.actions == list(want["actions"])
    assert np.array_equal(r.states, want["states"])
    assert (r.landed_successfully, r.crashed, r.game_over) == (want["landed_successfully"], want["crashed"], want["game_over"])
    assert r.penalty == want["penalty"] and (r.x, r.y, r.fuel) == (want["x"], want["y"], want["fuel"])


def test_trained_file_plays(mods, tmp_path):
    """configs[4] end to end in small: main_train_d (8_16_32_16_8) writes data/lunar_lander_last.txt, nn_play loads it unchanged"""
    import subprocess
    _, _, _, play = mods
    from test_gpu_host import BIN, write_config
    d = str(tmp_path)
    write_config(d, **{"NNConfig.hlayers": "8,16,32,16,8", "NNConfig.epochs": 3, "NNConfig.population_size": 64,
                       "NNTrainingCfg.layout_nb": 4})
    p = subprocess.run([os.path.join(BIN, "main_train_d")], cwd=d, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    cfg = os.path.join(d, "config.txt")
    r = play.nn_play(cfg, os.path.join(d, "data", "lunar_lander_last.txt"), seed=7, log=lambda s: None)
    assert r.steps > 0 and r.frames >= r.steps and (r.game_over or r.frames == 5000 + 180)
    assert ROOT


REF_SEQUENCE = r'''
# the literal call sequence of the reference: main.py:19-28 (imports by top-level name), mod_nn_play.py:26-59 (NeuralNetwork),
# main.py:127-142 (logic + pads), :192-238 (nn_play branch of the loop) — pygame removed, time = frame counter at 60 fps
import sys, os, json
import numpy as np
import lunar_lander_cpp
import cpp_nn_py
from lunar_lander_cpp import GameLogicCppDouble as GameLogicCpp
from lunar_lander_cpp import generate_random_pad_positions

class NeuralNetwork():
    def __init__(self):
        self._net = cpp_nn_py.ANN_MLP_GA_double()
        self._net.SetName("lunar_lander")
        full_path = os.path.join("./data/", "lunar_lander_last.txt")
        self._net.Deserialize(full_path)
        self._nGen = self._net.GetEpochs()
        self._nnsize = self._net.GetNetworkSize()
    def get_action(self, current_state):
        inputs = np.array(current_state, dtype=np.float64)
        return self._net.feedforwardIndex(inputs, 0)

logic = GameLogicCpp(False, 'config.txt')
spad_x1, lpad_x1 = generate_random_pad_positions(960, 80, 200, seed=int(sys.argv[1]), force_left_to_right=False, force_right_to_left=False)
logic.reset(spad_x1, lpad_x1)
NN = NeuralNetwork()
game_over = has_started = False
penalty, current_game_steps, frame = 0.0, 0, 0
state_buffer = np.zeros(5, dtype=np.float64)
actions = []
while frame < 5180 and not game_over:
    action = 0
    logic.get_state(state_buffer)
    nn_action = NN.get_action(state_buffer)
    if not has_started:
        if frame * 1000.0 / 60.0 > 2000:
            action = 1; has_started = True
        elif nn_action != 0:
            action = 1; has_started = True
    else:
        action = nn_action
    if has_started and not game_over:
        done = logic.update(action, state_buffer)
        penalty += logic.calculate_step_penalty(action)
        current_game_steps += 1
        actions.append(int(action))
        if done:
            game_over = True
    if game_over:
        penalty += logic.calculate_terminal_penalty(current_game_steps)
    frame += 1
print(json.dumps({"gen": NN._nGen, "size": list(NN._nnsize), "pads": [spad_x1, lpad_x1], "steps": current_game_steps, "frames": frame,
                  "landed": bool(logic.landed_successfully), "crashed": bool(logic.crashed), "penalty": penalty, "x": logic.x, "y": logic.y,
                  "fuel": logic.fuel, "actions": actions, "modules": [lunar_lander_cpp.__file__, cpp_nn_py.__file__]}))
'''


def test_reference_imports_and_call_sequence_in_a_fresh_interpreter(mods, tmp_path):
    """INTEGRATION.md §3 literally: a fresh interpreter whose only extra sys.path entry is lunar_lander_b200/pymodules imports
    `lunar_lander_cpp` and `cpp_nn_py` BY THOSE TOP-LEVEL NAMES (main.py:19-28, mod_nn_play.py:15-20), loads the
    data/lunar_lander_last.txt written by main_train_f and flies main.py's nn_play loop; the result equals the in-package
    headless loop on the same file and seed."""
    import json
    import subprocess
    import sys
    _, _, _, play = mods
    from test_gpu_host import BIN, write_config
    d = str(tmp_path)
    write_config(d, **{"NNConfig.hlayers": "8,16,8", "NNConfig.epochs": 3, "NNConfig.population_size": 64, "NNTrainingCfg.layout_nb": 4})
    p = subprocess.run([os.path.join(BIN, "main_train_f")], cwd=d, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    assert os.path.exists(os.path.join(d, "data", "lunar_lander_last.txt"))
    env = {k: v for k, v in os.environ.items() if k != "PYTHONPATH"}
    env["PYTHONPATH"] = play.PYMODULES
    q = subprocess.run([sys.executable, "-c", REF_SEQUENCE, "7"], cwd=d, capture_output=True, text=True, timeout=300, env=env)
    assert q.returncode == 0, q.stderr + q.stdout
    got = json.loads(q.stdout.strip().splitlines()[-1])
    assert all(os.path.dirname(m) == play.PYMODULES for m in got["modules"])  # the native extension modules, nothing else
    assert got["gen"] == 3 and got["size"] == [5, 8, 16, 8, 4]
    r = play.nn_play(os.path.join(d, "config.txt"), os.path.join(d, "data", "lunar_lander_last.txt"), seed=7, log=lambda s: None, fps=60.0)
    assert (got["pads"], got["steps"], got["frames"]) == ([r.spad_x1, r.lpad_x1], r.steps, r.frames)
    assert got["actions"] == r.actions and got["penalty"] == r.penalty
    assert (got["landed"], got["crashed"], got["x"], got["y"], got["fuel"]) == (r.landed_successfully, r.crashed, r.x, r.y, r.fuel)
