"""Multi-GPU step of the sweep: candidates are sharded (contiguous blocks, or round-robin
with plc_vl_upload_candidates_strided, which balances the ranks better), the terrain grid and cells are replicated, and the only exchange is one all-gather of 16 bytes per rank
— {float64 best_score, int64 best_global_index} — followed by the same deterministic argmax on
every rank (highest score, lowest global index on ties: the reference's first-maximum rule,
virtual_lidar.cpp:464-475).  `torch.distributed` is the plumbing (NCCL over NVLink on the GPU box,
gloo in the CPU tests)."""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def shard_range(n_candidates: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block [begin, end) of rank `rank`; blocks ascend with rank, so the lowest global
    index among equal scores lives on the lowest rank."""
    return rank * n_candidates // world, (rank + 1) * n_candidates // world


class _DevicePair:
    """__cuda_array_interface__ view of a context's 16-byte result pair as two int64 words."""

    def __init__(self, ptr: int):
        self.__cuda_array_interface__ = {"shape": (2,), "typestr": "<i8", "data": (ptr, False), "version": 3,
                                         "strides": None}


def result_pair_tensor(ctx) -> torch.Tensor:
    """Zero-copy torch view (int64[2] on the context's device) of {best_score bits, best_index}."""
    return torch.as_tensor(_DevicePair(ctx.result_device_ptr()), device=torch.device("cuda", ctx.device))


def pack_pair(score: float, index: int, device="cpu") -> torch.Tensor:
    bits = np.array([score], np.float64).view(np.int64)[0]
    return torch.tensor([int(bits), int(index)], dtype=torch.int64, device=device)


def pick_best(pairs: torch.Tensor) -> tuple[float, int]:
    """pairs: int64 (world, 2).  Deterministic first maximum over ranks."""
    host = pairs.detach().cpu().numpy()
    scores = host[:, 0].copy().view(np.float64)
    idx = host[:, 1]
    best_s, best_i = -np.inf, -1
    for s, i in zip(scores, idx):
        if i < 0:
            continue
        if s > best_s or (s == best_s and i < best_i):
            best_s, best_i = float(s), int(i)
    return best_s, best_i


def all_gather_best(local_pair: torch.Tensor, out: torch.Tensor | None = None, group=None) -> torch.Tensor:
    """One collective: gathers every rank's 16-byte pair.  Returns int64 (world, 2)."""
    world = dist.get_world_size(group)
    if out is None:
        out = torch.empty((world, 2), dtype=torch.int64, device=local_pair.device)
    dist.all_gather_into_tensor(out.view(-1), local_pair.contiguous(), group=group)
    return out
