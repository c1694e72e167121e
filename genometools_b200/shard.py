"""Multi-GPU sharding of the path: one process per GPU, shards = prefix-bucket code ranges.

This is the reference's `-parts` split (gt_suftabparts_new, src/match/sfx-partssuf.c:172-343) used
as a device map: rank r sorts and LCP-annotates part r; outputs are concatenated in rank order,
followed by the tail.  Nothing but the plan travels between ranks (the packed sequence is
broadcast once before the path starts), so the functions here only need a process group for
agreement checks and for gathering the per-shard statistics - they work on gloo (CPU tests) and
nccl alike.
"""
from __future__ import annotations

import numpy as np


def shard_plan(leftborder: np.ndarray, prefixlength: int, world: int):
    """[(mincode, maxcode, width, offset)] per rank; ranks beyond the number of non-empty parts get
    width 0 and an empty range marker (mincode > maxcode)."""
    from . import sfx
    parts = sfx.suftabparts(world, leftborder, prefixlength)
    plan = []
    for r in range(world):
        if r < len(parts):
            lo, hi, w = parts[r]
            plan.append((lo, hi, w, int(leftborder[lo])))
        else:
            plan.append((1, 0, 0, int(leftborder[-1])))
    return plan


def agree_on_plan(leftborder: np.ndarray | None, prefixlength: int, group=None):
    """Rank 0's bucket table decides; every rank ends up with the identical plan.  Returns
    (plan, leftborder)."""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    meta = torch.zeros(1, dtype=torch.int64)
    if rank == 0:
        meta[0] = leftborder.shape[0]
    dist.broadcast(meta, 0, group=group)
    t = torch.zeros(int(meta[0]), dtype=torch.int64)
    if rank == 0:
        t.copy_(torch.from_numpy(leftborder.astype(np.int64)))
    dist.broadcast(t, 0, group=group)
    lb = t.numpy().astype(np.uint32)
    return shard_plan(lb, prefixlength, world), lb


def gather_shard_stats(width: int, lcptabsum: float, maxbranchdepth: int, numlarge: int, longest: int, group=None):
    """All-gather of the per-part values the host needs for the .prj file; returns the combined
    (total_width, lcptabsum, maxbranchdepth, largelcpvalues, longest)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    mine = torch.tensor([float(width), float(lcptabsum), float(maxbranchdepth), float(numlarge), float(longest)],
                        dtype=torch.float64)
    allv = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(allv, mine, group=group)
    a = torch.stack(allv).numpy()
    longest_vals = a[:, 4][a[:, 4] >= 0]
    return (int(a[:, 0].sum()), float(a[:, 1].sum()), int(a[:, 2].max()), int(a[:, 3].sum()),
            int(longest_vals[0]) if longest_vals.size else -1)
