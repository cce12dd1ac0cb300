"""Object-level sharding over the GPUs of one box.

Objects are independent (the reference loops over them serially, src/LocalMapping_util.cc:213,382), so
each rank solves a contiguous block of the object list with no collective inside the GN loop; one
all-gather of the fixed 82-float records (t_cam_obj 16 | code 64 | is_good | loss) returns the
results in object order.  ``torch.distributed`` is plumbing only (NCCL on GPUs, gloo in CPU tests).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np

RECORD_LEN = 82


def partition(n_objects: int, world_size: int) -> List[Tuple[int, int]]:
    """Contiguous [lo, hi) blocks, sizes differing by at most one, earlier ranks get the extras."""
    base, extra = divmod(n_objects, world_size)
    out, lo = [], 0
    for r in range(world_size):
        hi = lo + base + (1 if r < extra else 0)
        out.append((lo, hi))
        lo = hi
    return out


def balanced_order(costs: Sequence[float], world_size: int) -> List[List[int]]:
    """Greedy longest-processing-time assignment by per-object cost (rays x samples): objects differ
    in work, so blocks of equal COUNT can be unbalanced.  Returns the object indices of each rank."""
    order = sorted(range(len(costs)), key=lambda i: -costs[i])
    load = [0.0] * world_size
    ranks: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        ranks[r].append(i)
        load[r] += costs[i]
    return [sorted(r) for r in ranks]


def gather_records(local_records, counts: Sequence[int], group=None):
    """All-gather per-rank record blocks (n_r, 82) into (sum n_r, 82) in rank order.  ``local_records``
    is a torch tensor on the rank's device (CUDA for NCCL, CPU for gloo).  Ragged counts are handled
    by padding to the maximum."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    mx = max(counts)
    pad = torch.zeros((mx, RECORD_LEN), dtype=torch.float32, device=local_records.device)
    pad[: local_records.shape[0]] = local_records
    out = torch.empty((world * mx, RECORD_LEN), dtype=torch.float32, device=local_records.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    parts = [out[r * mx: r * mx + counts[r]] for r in range(world)]
    return torch.cat(parts, 0)


def scatter_back(gathered: np.ndarray, rank_indices: Sequence[Sequence[int]]) -> np.ndarray:
    """Undo ``balanced_order``: rows of ``gathered`` are in rank order; return them in object order."""
    n = sum(len(r) for r in rank_indices)
    out = np.empty((n, gathered.shape[1]), gathered.dtype)
    k = 0
    for idx in rank_indices:
        for i in idx:
            out[i] = gathered[k]
            k += 1
    return out
