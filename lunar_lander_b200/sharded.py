"""Population sharded over the GPUs of one box: one process per GPU, contiguous global member ranges per rank
(SURVEY.md §8e).  Members are independent inside a generation (src/nn_engine_trainer.cpp:526-530), so evaluation
needs no data-path collective; a generation has exactly two small exchanges:

  1. all-gather of every rank's `top` best (fitness, global index) pairs       -> global elite indices
  2. all-reduce (integer SUM over bit patterns == broadcast from the owners) of the [top][P] elite matrix

Both go through ``torch.distributed`` (NCCL on GPUs, gloo in the CPU tests of the host logic).  The engine's
kernels and the collectives are ordered on one CUDA stream (the engine is switched to torch's current stream).
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np


def pad_candidates(fitness: np.ndarray, index: np.ndarray, top: int) -> np.ndarray:
    """[top, 2] float64 block a rank contributes to the all-gather; short shards pad with (+inf, 2^53)."""
    out = np.empty((top, 2), dtype=np.float64)
    out[:, 0] = np.inf
    out[:, 1] = float(1 << 53)
    k = min(len(fitness), top)
    out[:k, 0] = fitness[:k]
    out[:k, 1] = index[:k].astype(np.float64)  # exact: indices < 2^32
    return out


def merge_candidates(blocks: np.ndarray, top: int) -> Tuple[np.ndarray, np.ndarray]:
    """Global elites from the gathered [world * top, 2] blocks: ascending (fitness, index), -0.0 ties with +0.0 —
    the order the device top-k uses on a single shard, so the result does not depend on the GPU count."""
    f = blocks[:, 0] + 0.0  # folds -0.0 onto +0.0
    i = blocks[:, 1]
    order = np.lexsort((i, f))[:top]
    return f[order], i[order].astype(np.uint64)


class ShardedGeneration:
    """evaluate -> select -> breed for one rank's shard, with the two exchanges above."""

    def __init__(self, engine, group=None, device=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.eng, self.group = engine, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        n64 = engine.elite_buffer_bytes() // 8
        self.elite_buf = torch.zeros(n64, dtype=torch.int64, device=self.device)
        self.cand_local = torch.empty((engine.top, 2), dtype=torch.float64, device=self.device)
        self.cand_all = torch.empty((self.world * engine.top, 2), dtype=torch.float64, device=self.device)
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def use_current_stream(self):
        self.eng.set_stream(self.torch.cuda.current_stream().cuda_stream)

    def select_and_breed(self, elitism: bool = True) -> np.ndarray:
        torch, dist, eng = self.torch, self.dist, self.eng
        f, idx = eng.local_topk()                       # K2 on this shard (+ top x 16 B to the host)
        self.d2h_bytes += 16 * len(f)
        block = pad_candidates(f, idx, eng.top)
        self.cand_local.copy_(torch.from_numpy(block), non_blocking=False)
        self.h2d_bytes += block.nbytes
        if self.world > 1:
            dist.all_gather_into_tensor(self.cand_all, self.cand_local, group=self.group)
        else:
            self.cand_all.copy_(self.cand_local)
        blocks = self.cand_all.cpu().numpy()
        self.d2h_bytes += blocks.nbytes
        _, elite_idx = merge_candidates(blocks, eng.top)
        eng.gather_elites(elite_idx, self.elite_buf.data_ptr())  # K3: owned rows, zeros elsewhere
        self.h2d_bytes += 8 * len(elite_idx)
        if self.world > 1:
            dist.all_reduce(self.elite_buf, op=dist.ReduceOp.SUM, group=self.group)
        eng.set_elites(self.elite_buf.data_ptr())
        eng.create_population(elitism)                   # K4 on this shard
        eng.update_epochs(1)
        return elite_idx

    def step(self, max_steps: int, elitism: bool = True) -> np.ndarray:
        self.eng.evaluate_device(max_steps)              # K1 (+ shard statistics)
        return self.select_and_breed(elitism)
