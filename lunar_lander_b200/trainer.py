"""Training::NNEngineTrainer (src/nn_engine_trainer.h:31-40) for a population sharded over the GPUs of one box.

Same public interface as the reference's trainer and as the C++ mirror in ``host/`` — ``init() / load(step) / save() / train()``,
same config.txt keys, log lines, file names, stagnation / layout-reset schedule — with one process per GPU
(``torchrun --nproc-per-node N -m lunar_lander_b200.trainer``): every rank evaluates its contiguous member range on all
layouts, the elites are chosen from an all-gather of each rank's top candidates and broadcast as one all-reduce
(``sharded.ShardedGeneration``), rank 0 prints and writes the checkpoints.  With one process it is the single-GPU trainer with
device-side selection (``GpuCfg.select = device`` of the C++ mirror), and prints the same log.

The environment, the forward pass, the GA: all on the GPU behind include/ll_gpu.h; nothing here computes them.
"""
from __future__ import annotations

import os
import sys
import time
from typing import List, Optional

import numpy as np

from . import _abi
from .config import Config
from .engine import Engine, make_layouts, peek_weight_file
from .sharded import ShardedGeneration


def _mmss(seconds: float) -> str:
    s = int(seconds)
    return "00:00" if s < 0 else f"{s // 60:02d}:{s % 60:02d}"


class NNEngineTrainer:
    def __init__(self, cfg: Config, precision: int = 64, mode: Optional[int] = None, out=sys.stdout):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.cfg, self.precision, self.out = cfg, precision, out
        self.mode = mode if mode is not None else (_abi.LL_MODE_EXACT if precision == 64 else _abi.LL_MODE_FAST)
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        g = cfg
        # the values the reference's constructor caches (src/nn_engine_trainer.cpp:29-65)
        self.name, self.save_dir = g.getString("NNConfig.name"), g.getString("NNConfig.save_path_nn")
        self.save_enabled, self.overwrite = g.getBool("NNConfig.save_nn"), g.getBool("NNConfig.overwrite")
        self.save_interval, self.epochs = g.getInt("NNConfig.save_interval"), g.getInt("NNConfig.epochs")
        self.layout_nb, self.ltr_ratio = g.getInt("NNTrainingCfg.layout_nb"), g.getDouble("NNTrainingCfg.left_right_ratio")
        self.seed, self.elitism = g.getInt("NNConfig.seed"), g.getBool("NNConfig.elitism")
        self.injection = g.getDouble("NNConfig.random_injection_ratio")
        self.sizes: List[int] = g.layer_sizes()
        self.max_steps, self.reset_period = g.getInt("GameCfg.max_steps"), g.getInt("NNTrainingCfg.reset_period")
        self.lr_unit = int(np.float32(g.getDouble("NNTrainingCfg.tp_landed_lr_bonus")) if precision == 32 else
                           g.getDouble("NNTrainingCfg.tp_landed_lr_bonus"))  # an int in the reference (src/nn_engine_trainer.h:98)
        self.stagnated, self.last_best = 0, float(np.finfo(np.float64).max)
        self.generation = 0
        self.eng: Optional[Engine] = None
        self.sg: Optional[ShardedGeneration] = None
        self.directions = None  # layout_directions outlives a layout reset (balance_layouts' early return, :214)
        self.history: List[dict] = []
        self.t0 = time.monotonic()

    def _say(self, text: str = "") -> None:
        if self.rank == 0:
            print(text, file=self.out, flush=True)

    def _engine(self, sizes, population, top, activation) -> Engine:
        if self.eng is not None:
            self.eng.close()
        device = self.torch.cuda.current_device()
        eng = Engine(self.cfg, sizes=sizes, population=population, top=top, precision=self.precision, mode=self.mode,
                     rank=self.rank, world=self.world, device=device, activation=activation)
        self.eng, self.sg = eng, ShardedGeneration(eng)
        self.sg.use_current_stream()
        return eng

    # ---- init / load / save: src/nn_engine_trainer.cpp:67-183
    def init(self) -> bool:
        try:
            if self.save_enabled and self.save_dir and self.rank == 0:
                os.makedirs(self.save_dir, exist_ok=True)
            eng = self._engine(self.sizes, self.cfg.getInt("NNConfig.population_size"), self.cfg.getInt("NNConfig.top_individuals"),
                               self.cfg.getInt("NNConfig.activation_id"))
            eng.set_population_strategy(_abi.LL_STRATEGY_MIXED_WITH_RANDOM_INJECTION, self.injection)
            eng.create_population(self.elitism)
        except (OSError, RuntimeError) as e:
            print(f"Error initializing the GPU network: {e}", file=sys.stderr)
            self.eng = None
            return False
        self.generation, self.t0 = 0, time.monotonic()
        return True

    def load(self, step: str = "last") -> bool:
        path = os.path.join(self.save_dir, f"{self.name}_{step}.txt")
        if not os.path.exists(path):
            print(f"Error loading network: File not found at {path}", file=sys.stderr)
            return False
        try:
            h = peek_weight_file(path)
            eng = self._engine(h["sizes"], h["population"], h["top"], h["act"])
            eng.deserialize(path)  # every rank reads the file: the parent set is replicated, the shard's children are rebuilt
            self.generation, self.sizes, self.t0 = eng.get_epochs(), list(h["sizes"]), time.monotonic()
            return True
        except RuntimeError as e:
            print(f"Error loading network from {path}: {e}", file=sys.stderr)
            return False

    def save(self) -> bool:
        if self.eng is None:
            print("Save error: Network not initialized.", file=sys.stderr)
            return False
        if self.rank != 0:
            return True
        if self.overwrite:
            fname = f"{self.name}_latest_overwrite.txt"
        else:
            fname = f"{self.name}_{'_'.join(str(s) for s in self.sizes)}_s_{self.generation:04d}.txt"
        numbered, last = os.path.join(self.save_dir, fname), os.path.join(self.save_dir, f"{self.name}_last.txt")
        try:
            self.eng.serialize(numbered, self.name)
            with open(numbered, "rb") as src, open(last, "wb") as dst:
                dst.write(src.read())
            return True
        except (OSError, RuntimeError) as e:
            print(f"Error during network serialization/copy: {e}", file=sys.stderr)
            return False

    # ---- one generation over all ranks
    def _draw_layouts(self, seed_base: int):
        spad, lpad, d, sc, lc = make_layouts(self.cfg, seed_base, self.layout_nb, self.directions)
        self.directions = d
        self.eng.set_layouts(spad, lpad, d)
        return int(np.sum(sc < lc))

    def _generation(self) -> dict:
        torch, dist, eng = self.torch, self.dist, self.eng
        eng.evaluate_device(self.max_steps)
        _, _, _, st = eng.read_results(want_arrays=False)
        mine = torch.tensor([st.best_fitness, float(st.best_index), st.best_lr, st.sum_fitness, st.sum_steps], dtype=torch.float64,
                            device="cuda")
        if self.world > 1:
            every = torch.empty((self.world, 5), dtype=torch.float64, device="cuda")
            dist.all_gather_into_tensor(every, mine)
        else:
            every = mine[None, :]
        every = every.cpu().numpy()
        order = np.lexsort((every[:, 1], every[:, 0] + 0.0))  # (fitness, index), -0.0 ties with +0.0
        best = every[order[0]]
        elite_idx = self.sg.select_and_breed(self.elitism)
        assert int(elite_idx[0]) == int(best[1])
        self.generation = eng.get_epochs()
        return dict(generation=self.generation, best_member=int(best[1]), best_fitness=float(best[0]), best_lr=float(best[2]),
                    sum_fitness=float(every[:, 3].sum()), sum_steps=float(every[:, 4].sum()))

    def train(self) -> None:
        if self.eng is None:
            print("Train error: Network not initialized or loaded.", file=sys.stderr)
            return
        eng = self.eng
        self._say(f"Starting C++ training for {self.epochs} generations...")
        self._say(f"Population size: {eng.pop_global}, Top performers: {eng.top}, Elitism: {int(self.elitism)}")
        self._say("Network structure: " + "-".join(str(s) for s in self.sizes))
        ltr = self._draw_layouts(self.seed)
        self._say(f"Layouts after balancing: {ltr} left-to-right, {self.layout_nb - ltr} right-to-left.")
        per_layout, per_member = float(self.layout_nb), float(eng.pop_global)
        for k in range(self.epochs):
            g0 = time.monotonic()
            shown = self.generation + 1
            if self.stagnated >= self.reset_period:
                self._say(f"--- Stagnation detected. Resetting training layouts after {self.stagnated} generations. --- ")
                self._draw_layouts(self.seed + self.generation + k)  # src/nn_engine_trainer.cpp:344-345
                self.stagnated, self.last_best = 0, float(np.finfo(np.float64).max)
            self._say(f"\n--- Generation {shown} ---")
            rep = self._generation()
            if rep["best_member"] == 0 and rep["best_fitness"] >= self.last_best - 1e-5:  # :382-402
                self.stagnated += 1
            else:
                self.stagnated = 0
            self.last_best = rep["best_fitness"]
            g1 = time.monotonic()
            rep["seconds"] = g1 - g0
            self.history.append(rep)
            landings = 0 if self.lr_unit == 0 else int(round(rep["best_lr"] / self.lr_unit))
            self._say(f"Generation {self.generation} complete. (Took: {_mmss(g1 - g0)}, Total: {_mmss(g1 - self.t0)})")
            self._say(f"  Best Avg Fitness: {rep['best_fitness'] / per_layout:.4f} (Member {rep['best_member']})")
            self._say(f"  Min Number of Landing in Both Directions: {landings}")
            self._say(f"  Avg Fitness:      {rep['sum_fitness'] / per_member / per_layout:.4f}")
            self._say(f"  Avg Steps/Layout: {rep['sum_steps'] / per_member / per_layout:.1f}")
            if self.save_enabled and (self.generation % self.save_interval == 0 or k == self.epochs - 1):
                self.save()
        self._say("\nC++ Training finished.")


def main(argv=None) -> int:
    """``torchrun --nproc-per-node N -m lunar_lander_b200.trainer [--float] [--continue] [--step N] [--config F] [--set k=v]``"""
    import argparse

    import torch
    import torch.distributed as dist
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="config.txt")
    ap.add_argument("--set", action="append", default=[], metavar="Section.key=value")
    ap.add_argument("--float", action="store_true", help="main_train_f (fp32; GpuCfg.mode = fast unless set) instead of main_train_d")
    ap.add_argument("--continue", dest="resume", action="store_true")
    ap.add_argument("--step", type=int, default=None)
    a = ap.parse_args(argv)
    if not torch.cuda.is_available():
        print("FATAL: no CUDA device (this trainer has no CPU fallback)", file=sys.stderr)
        return 1
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    cfg = Config()
    if not cfg.loadConfiguration(a.config):
        print(f"FATAL: Could not load configuration from {a.config}. Exiting.", file=sys.stderr)
        return 1
    for kv in a.set:
        k, _, v = kv.partition("=")
        cfg.set(k, v)
    precision = 32 if a.float else 64
    mode_name = cfg.kv.get("GpuCfg.mode", "fast" if a.float else "exact")
    tr = NNEngineTrainer(cfg, precision, _abi.LL_MODE_FAST if mode_name == "fast" else _abi.LL_MODE_EXACT)
    tr._say(f"Using data type: {'float' if a.float else 'double'}")
    if not tr.init():
        print("Failed to initialize NNEngineTrainer.", file=sys.stderr)
        return 1
    if a.resume or a.step is not None:
        ckpt = "last" if a.step is None else f"gen_{a.step:04d}"
        tr._say(f"Attempting to load network from checkpoint: {ckpt}")
        if tr.load(ckpt):
            tr._say(f"Successfully loaded network from checkpoint '{ckpt}'.")
        else:
            print(f"Failed to load network from checkpoint '{ckpt}'. Training will start from scratch (using the newly initialized "
                  "network).", file=sys.stderr)
            tr.init()
    else:
        tr._say("Starting training from scratch (no --continue specified, or --step without --continue).")
    tr.train()
    tr._say("Training process finished in C++.")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
