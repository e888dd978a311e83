"""config.txt reader with the interface of ma-libs' ``Config`` namespace as the reference uses it
(Config::loadConfiguration / getDouble / getInt / getBool / getString / getVector*; call sites
src/main_train.cpp:17, src/game_logic.cpp:42-95, src/nn_engine_trainer.cpp:29-65).

Syntax (config.txt): ``Section.key = value   # comment``; vectors are comma separated; booleans ``true``/``false``;
unknown keys are kept and ignored (older sample configs carry extra keys).
"""
from __future__ import annotations

import re
from typing import Dict, List

from . import _abi


class Config:
    """Instance-based (the reference's store is process-global; a module-level default mirrors that)."""

    def __init__(self, path: str | None = None, overrides: Dict[str, object] | None = None):
        self.kv: Dict[str, str] = {}
        if path:
            if not self.loadConfiguration(path):
                raise RuntimeError(f"FATAL: Could not load configuration from {path}. Exiting.")  # game_logic.cpp:42-45
        for k, v in (overrides or {}).items():
            self.set(k, v)

    def loadConfiguration(self, path: str) -> bool:
        try:
            with open(path, "r") as f:
                lines = f.readlines()
        except OSError:
            return False
        for line in lines:
            line = line.split("#", 1)[0]
            if "=" not in line:
                continue
            k, v = line.split("=", 1)
            if k.strip():
                self.kv[k.strip()] = v.strip()
        return True

    def set(self, key: str, value) -> None:
        if isinstance(value, bool):
            value = "true" if value else "false"
        elif isinstance(value, (list, tuple)):
            value = ",".join(str(x) for x in value)
        self.kv[key] = str(value)

    def _raw(self, key: str) -> str:
        if key not in self.kv:
            raise KeyError(f"Config: missing key {key}")
        return self.kv[key]

    def getString(self, key: str) -> str:
        return self._raw(key)

    def getDouble(self, key: str) -> float:
        return float(self._raw(key))

    def getInt(self, key: str) -> int:
        m = re.match(r"\s*[+-]?\d+", self._raw(key))  # strtol semantics: leading integer, 0 if none
        return int(m.group(0)) if m else 0

    def getBool(self, key: str) -> bool:
        return self._raw(key) in ("true", "True", "TRUE", "1")

    def _split(self, key: str) -> List[str]:
        return [s.strip() for s in self._raw(key).split(",") if s.strip()]

    def getVectorDouble(self, key: str) -> List[float]:
        return [float(s) for s in self._split(key)]

    getVectorFloat = getVectorDouble

    def getVectorInt(self, key: str) -> List[int]:
        return [int(s) for s in self._split(key)]

    # ---- what the C ABI needs -------------------------------------------------------------------------------------
    def env_cfg(self) -> _abi.EnvCfg:
        """Every value GameLogicCpp<T>'s constructor reads (src/game_logic.cpp:48-95), as doubles."""
        e = _abi.EnvCfg()
        e.pad_y1 = self.getDouble("GameCfg.pad_y1")
        e.terrain_y = self.getDouble("GameCfg.terrain_y")
        e.max_vx = self.getDouble("GameCfg.max_vx")
        e.max_vy = self.getDouble("GameCfg.max_vy")
        e.g = self.getDouble("PlanetCfg.g")
        e.mu_x = self.getDouble("PlanetCfg.mu_x")
        e.mu_y = self.getDouble("PlanetCfg.mu_y")
        e.lander_width = self.getDouble("LanderCfg.width")
        e.lander_height = self.getDouble("LanderCfg.height")
        e.max_fuel = float(self.getInt("LanderCfg.max_fuel"))
        e.lpad_x1 = self.getDouble("GameCfg.lpad_x1")
        e.lpad_width = self.getDouble("GameCfg.lpad_width")
        e.display_width = self.getDouble("DisplayCfg.width")
        e.display_height = self.getDouble("DisplayCfg.height")
        e.spad_width = self.getDouble("GameCfg.spad_width")
        e.x0, e.y0 = self.getVectorDouble("GameCfg.x0")[:2]
        e.vx0, e.vy0 = self.getVectorDouble("GameCfg.v0")[:2]
        e.ax0, e.ay0 = self.getVectorDouble("GameCfg.a0")[:2]
        for f, k in (("sp_dist_factor", "sp_dist_factor"), ("sp_action_factor", "sp_action_factor"),
                     ("tp_steps_factor", "tp_steps_factor"), ("tp_dist_factor", "tp_dist_factor"),
                     ("tp_landed_bonus", "tp_landed_bonus"), ("tp_landed_lr_bonus", "tp_landed_lr_bonus"),
                     ("tp_fuel_bonus_factor", "tp_fuel_bonus_factor"), ("tp_crashed_penalty", "tp_crashed_penalty"),
                     ("tp_crash_v_mag_factor", "tp_crash_v_mag_factor"), ("tp_no_fuel_penalty", "tp_no_fuel_penalty")):
            setattr(e, f, self.getDouble("NNTrainingCfg." + k))
        return e

    def layer_sizes(self) -> List[int]:
        """{5, hlayers..., 4}: src/nn_engine_trainer.cpp:45-48."""
        return [5] + self.getVectorInt("NNConfig.hlayers") + [4]


def default_config_path() -> str:
    import os
    return os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "config.txt")
