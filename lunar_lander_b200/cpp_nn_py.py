"""Drop-in for ma-libs' Python module ``cpp_nn_py`` as the reference uses it (mod_nn_play.py:19-56): ``ANN_MLP_GA_double`` /
``ANN_MLP_GA_float`` with ``SetName``, ``Deserialize``, ``Serialize``, ``GetEpochs``, ``GetNetworkSize``, ``GetPopSize``,
``GetTopPerformersSize`` and ``feedforwardIndex(inputs, member)``.  The network lives on the GPU behind include/ll_gpu.h;
``feedforwardIndex`` is one ``ll_feedforward_index`` call (the exact-arithmetic forward pass of the episode kernel).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional

import numpy as np

from . import _abi
from ._abi import check, lib
from .config import Config, default_config_path
from .engine import Engine, peek_weight_file


class _Net:
    _precision, _dtype = 64, np.float64

    def __init__(self, config: Optional[Config] = None):
        self._cfg = config
        self._eng: Optional[Engine] = None
        self._name = "net"

    def _config(self) -> Config:
        if self._cfg is None:
            path = "config.txt" if os.path.exists("config.txt") else default_config_path()
            self._cfg = Config(path)
        return self._cfg

    def _need(self) -> Engine:
        if self._eng is None:
            raise RuntimeError("ANN_MLP_GA: the network has no shape yet (Deserialize first)")
        return self._eng

    def SetName(self, name: str) -> None:
        self._name = str(name)

    def Deserialize(self, path: str) -> None:
        """The file decides sizes / population / top / activation (an empty net takes its shape from it,
        src/nn_engine_trainer.cpp:104, 126-128)."""
        h = peek_weight_file(path)  # raises LLError (a RuntimeError) on a missing or malformed file
        if self._eng is not None:
            self._eng.close()
        self._eng = Engine(self._config(), sizes=h["sizes"], population=h["population"], top=h["top"], precision=self._precision,
                           mode=_abi.LL_MODE_EXACT, activation=h["act"])
        self._eng.deserialize(path)

    def Serialize(self, path: str) -> None:
        self._need().serialize(path, self._name)

    def GetEpochs(self) -> int:
        return self._need().get_epochs() if self._eng is not None else 0

    def GetNetworkSize(self) -> List[int]:
        return list(self._need().sizes)

    def GetPopSize(self) -> int:
        return self._need().pop_global

    def GetTopPerformersSize(self) -> int:
        return self._need().top

    def feedforwardIndex(self, inputs, member: int = 0) -> int:
        x = np.ascontiguousarray(inputs, dtype=self._dtype)
        if x.ndim != 1 or x.shape[0] < 5:
            raise RuntimeError("feedforwardIndex: inputs must be a 1-dimensional array of at least 5 values")
        return self._need().feedforward_index(x, int(member))


class ANN_MLP_GA_double(_Net):
    _precision, _dtype = 64, np.float64


class ANN_MLP_GA_float(_Net):
    _precision, _dtype = 32, np.float32
