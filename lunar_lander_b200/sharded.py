"""Population sharded over the GPUs of one box: one process per GPU, contiguous global member ranges per rank
(SURVEY.md §8e).  Members are independent inside a generation (src/nn_engine_trainer.cpp:526-530), so evaluation
needs no data-path collective; a generation has exactly two small exchanges:

  1. all-gather of every rank's `top` best (fitness, global index) pairs       -> global elite indices
  2. all-reduce (integer SUM over bit patterns == broadcast from the owners) of the [top][P] elite matrix

Both live BEHIND the C ABI (csrc/ll_comm.cu): the context owns the NCCL communicator, the candidate merge runs on the device and
the whole generation (K1 -> K2 -> all-gather -> merge -> K3 -> all-reduce -> K4) is enqueued on the context's stream without a host
round trip.  ``torch.distributed`` is used for one thing only: handing rank 0's 128-byte communicator id to the other processes.
``pad_candidates`` / ``merge_candidates`` are the host statement of the merge order (what the device merge must reproduce); the CPU
tests check them over gloo.
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
    """evaluate -> select -> breed for one rank's shard; the exchanges are the context's (ll_comm_init / ll_step_generation)."""

    def __init__(self, engine, group=None, device=None):
        import torch.distributed as dist
        self.eng, self.group = engine, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        if self.world > 1:
            box = [engine.comm_unique_id() if dist.get_rank(group) == 0 else None]
            dist.broadcast_object_list(box, src=0, group=group)
            engine.comm_init(box[0])

    def use_current_stream(self):
        import torch
        self.eng.set_stream(torch.cuda.current_stream().cuda_stream)

    def select_and_breed(self, elitism: bool = True) -> np.ndarray:
        elite_idx = self.eng.select(None)   # K2 -> NCCL all-gather -> device merge -> K3 -> NCCL all-reduce (+ top x 8 B to the host)
        self.d2h_bytes += 8 * len(elite_idx)
        self.eng.create_population(elitism)  # K4 on this shard
        self.eng.update_epochs(1)
        return elite_idx

    def step(self, max_steps: int, elitism: bool = True, want_stats: bool = False):
        """one generation, nothing but (optionally) the statistics crosses PCIe"""
        return self.eng.step_generation(max_steps, elitism, want_stats=want_stats)
