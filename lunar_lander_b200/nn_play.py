"""Headless ``main.py --mode=nn_play``: the reference's game loop (main.py:104-320) without pygame — no window, sprites or
sound (unavailable on this box); time is a frame counter at ``DisplayCfg.fps`` instead of ``pygame.time.get_ticks()``.

What is kept, line for line in behaviour (main.py:127-238):
  * pads from ``generate_random_pad_positions(width, spad_width, lpad_width, seed, force_ltr, force_rtl)``; ``logic.reset(spad, lpad)``;
  * the network is ``<NNConfig.name>_last.txt`` under ``NNConfig.save_path_nn`` loaded the way mod_nn_play.NeuralNetwork does
    (always the double build, member 0);
  * start gating: until the flight has started the lander is not updated; the first non-zero network action — or 2 s of
    frames — starts it with a forced UP (action 1); afterwards the network's action is used as is;
  * per updated frame ``penalty += calculate_step_penalty(action)``, at game over ``penalty += calculate_terminal_penalty(steps)``.

Everything numeric runs on the GPU: the lander through the pybind11 module ``lunar_lander_cpp`` (ll_sim_*), the policy through
the pybind11 module ``cpp_nn_py`` (ll_feedforward_index).  Both are imported by their top-level names from ``pymodules/`` — the
directory a user of the reference puts on ``sys.path`` in place of ``externals/ma-libs/build/Release`` (main.py:19-21,
mod_nn_play.py:15-17).
"""
from __future__ import annotations

import os
import sys
from dataclasses import dataclass, field
from typing import Callable, List, Optional

import numpy as np

from .config import Config

PYMODULES = os.path.join(os.path.dirname(os.path.abspath(__file__)), "pymodules")
if PYMODULES not in sys.path:
    sys.path.append(PYMODULES)
try:
    import cpp_nn_py  # noqa: E402
    import lunar_lander_cpp  # noqa: E402
except ModuleNotFoundError as e:  # no fallback: the native modules are the only implementation
    raise ImportError(f"{e}; build them with `make -C lunar_lander_b200/host` (or __graft_entry__.build())") from e


class NeuralNetwork:
    """mod_nn_play.NeuralNetwork (mod_nn_play.py:25-59)."""

    def __init__(self, cfg: Config, weights: Optional[str] = None):
        self._net = cpp_nn_py.ANN_MLP_GA_double()
        name = cfg.getString("NNConfig.name")
        self._net.SetName(name)
        path = weights or os.path.join(cfg.getString("NNConfig.save_path_nn"), f"{name}_last.txt")
        self._net.Deserialize(path)
        self._nGen = self._net.GetEpochs()
        self._nnsize = self._net.GetNetworkSize()

    @property
    def generation(self) -> int:
        return self._nGen

    @property
    def size(self) -> List[int]:
        return self._nnsize

    def get_action(self, current_state: np.ndarray) -> int:
        inputs = np.array(current_state, dtype=np.float64)
        try:
            return self._net.feedforwardIndex(inputs, 0)  # member 0 = the best of the last generation
        except Exception as e:  # the reference swallows the error and returns NOOP (mod_nn_play.py:57-59)
            print(f"Error during feedforward: {e}")
            return 0


@dataclass
class PlayResult:
    spad_x1: int
    lpad_x1: int
    frames: int                 # frames of the main loop, waiting ones included
    steps: int                  # logic updates (current_game_steps)
    forced_start_frame: int     # frame at which the flight started
    landed_successfully: bool
    crashed: bool
    game_over: bool
    penalty: float
    x: float
    y: float
    fuel: float
    actions: List[int] = field(default_factory=list)
    states: Optional[np.ndarray] = None   # [steps][5] post-update get_state(), when record=True


def nn_play(config_file: str = "config.txt", weights: Optional[str] = None, seed: Optional[int] = None,
            force_left_to_right: bool = False, force_right_to_left: bool = False, max_frames: Optional[int] = None,
            record: bool = False, log: Callable[[str], None] = print, fps: Optional[float] = None) -> PlayResult:
    """One nn_play episode.  ``max_frames`` bounds a network that never lands (the UI runs until the window is closed)."""
    cfg = Config(config_file)
    logic = lunar_lander_cpp.GameLogicCppDouble(False, config_file)
    width = cfg.getDouble("DisplayCfg.width")
    spad_w, lpad_w = cfg.getDouble("GameCfg.spad_width"), cfg.getDouble("GameCfg.lpad_width")
    spad_x1, lpad_x1 = lunar_lander_cpp.generate_random_pad_positions(width, int(spad_w), int(lpad_w), seed=seed,
                                                                      force_left_to_right=force_left_to_right,
                                                                      force_right_to_left=force_right_to_left)
    logic.reset(spad_x1, lpad_x1)
    nn = NeuralNetwork(cfg, weights)
    fps = float(fps if fps is not None else (cfg.getDouble("DisplayCfg.fps") if "DisplayCfg.fps" in cfg.kv else 60.0))
    if max_frames is None:
        max_frames = cfg.getInt("GameCfg.max_steps") + int(3 * fps)

    state = np.zeros(5, dtype=np.float64)
    has_started = game_over = False
    penalty, steps, frame, start_frame = 0.0, 0, 0, -1
    actions: List[int] = []
    states: List[np.ndarray] = []
    while frame < max_frames and not game_over:
        action = 0
        logic.get_state(state)
        nn_action = nn.get_action(state)
        if not has_started:
            elapsed_ms = frame * 1000.0 / fps  # pygame.time.get_ticks() - start_time
            if elapsed_ms > 2000:
                log("GameLogic: Forcing start (UP) after 2 seconds.")
                action, has_started = 1, True
            elif nn_action != 0:
                action, has_started = 1, True
            if has_started:
                start_frame = frame
        else:
            action = nn_action
        if has_started:
            done = logic.update(action, state)
            penalty += logic.calculate_step_penalty(action)
            steps += 1
            actions.append(action)
            if record:
                states.append(state.copy())
            if done:
                game_over = True
                log("Landing Successful!" if logic.landed_successfully else "Crashed!")
        if game_over:
            penalty += logic.calculate_terminal_penalty(steps)
        frame += 1
    return PlayResult(spad_x1, lpad_x1, frame, steps, start_frame, bool(logic.landed_successfully), bool(logic.crashed),
                      game_over, penalty, logic.x, logic.y, logic.fuel, actions, np.array(states) if record else None)


def main(argv=None) -> int:
    import argparse
    ap = argparse.ArgumentParser(description="headless nn_play (main.py --mode=nn_play without the window)")
    ap.add_argument("--config", default="config.txt")
    ap.add_argument("--weights", default=None, help="default: <NNConfig.save_path_nn>/<NNConfig.name>_last.txt")
    ap.add_argument("--seed", type=int, default=None)
    ap.add_argument("--left_to_right", action="store_true")
    ap.add_argument("--right_to_left", action="store_true")
    ap.add_argument("--episodes", type=int, default=1)
    a = ap.parse_args(argv)
    landed = 0
    for i in range(a.episodes):
        r = nn_play(a.config, a.weights, None if a.seed is None else a.seed + i, a.left_to_right, a.right_to_left)
        landed += int(r.landed_successfully)
        print(f"episode {i}: pads {r.spad_x1}->{r.lpad_x1} steps {r.steps} landed {r.landed_successfully} "
              f"fuel {r.fuel:.1f} penalty {r.penalty:.4f}")
    print(f"{landed}/{a.episodes} landed")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
