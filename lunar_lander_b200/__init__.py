"""lunar_lander_b200 — B200 (sm_100a) implementation of lunar-lander's neuro-evolution hot path.

The package holds only what that path needs: ``csrc/`` (CUDA kernels + the C ABI of include/ll_gpu.h, built into
``libll_gpu.so`` next to this file) and the host-side mirrors of the reference's interfaces for it: ``config.Config``,
``engine.Engine`` (the network + population behind the C ABI), ``trainer.NNEngineTrainer`` (the sharded multi-GPU trainer),
``nn_play`` (the headless nn_play loop over the pybind11 modules ``lunar_lander_cpp`` / ``cpp_nn_py`` built into ``pymodules/``), and in ``host/`` the C++ mirror (``main_train_d/_f``).
Importing it requires the built library; there is no CPU fallback.
"""
from . import _abi  # noqa: F401  (raises ImportError when libll_gpu.so has not been built)
from ._abi import LL_MODE_EXACT, LL_MODE_FAST, LLError  # noqa: F401
from .config import Config, default_config_path  # noqa: F401
from .engine import Engine, generate_random_pad_positions, make_layouts, peek_weight_file  # noqa: F401

__all__ = ["Config", "Engine", "LLError", "LL_MODE_EXACT", "LL_MODE_FAST", "default_config_path",
           "generate_random_pad_positions", "make_layouts", "peek_weight_file"]
