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


def balanced_equal_counts(costs: Sequence[float], world_size: int) -> List[List[int]]:
    """Like ``balanced_order`` but every rank gets the same NUMBER of objects (sizes differ by at most one): objects
    sorted by cost are dealt out in a snake (0..R-1, R-1..0, ...), which keeps the per-rank cost sums within one
    object's cost of each other.  Used for weak scaling, where the count per GPU is fixed."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    ranks: List[List[int]] = [[] for _ in range(world_size)]
    for k, i in enumerate(order):
        lap, pos = divmod(k, world_size)
        ranks[pos if lap % 2 == 0 else world_size - 1 - pos].append(i)
    return [sorted(r) for r in ranks]


def refine_equal_counts(costs: Sequence[float], shares: List[List[int]], max_swaps: int = 1000) -> List[List[int]]:
    """Local search on an equal-count dealing: while it narrows the spread, swap one object of the heaviest rank with
    one of the lightest (the pair whose cost difference is closest to half the gap).  The snake of
    ``balanced_equal_counts`` leaves the rank sums within one object's cost of each other - 3 % of a 32-object share;
    a few swaps bring them within a fraction of a percent.  Deterministic (ties by index), so every rank computes the
    same dealing from the same costs."""
    shares = [list(r) for r in shares]
    sums = [sum(costs[i] for i in r) for r in shares]
    for _ in range(max_swaps):
        hi = max(range(len(shares)), key=lambda r: (sums[r], -r))
        lo = min(range(len(shares)), key=lambda r: (sums[r], r))
        gap = sums[hi] - sums[lo]
        if hi == lo or gap <= 0:
            break
        best, best_err = None, gap                         # after a swap of difference d the gap becomes |gap - 2 d|
        for a in shares[hi]:
            for b in shares[lo]:
                d = costs[a] - costs[b]
                if d > 0 and abs(gap - 2 * d) < best_err:
                    best, best_err = (a, b), abs(gap - 2 * d)
        if best is None:
            break
        a, b = best
        shares[hi][shares[hi].index(a)] = b
        shares[lo][shares[lo].index(b)] = a
        sums[hi] -= costs[a] - costs[b]
        sums[lo] += costs[a] - costs[b]
    return [sorted(r) for r in shares]


def inball_cost(detection: dict, num_depth_samples: int) -> float:
    """Host-side proxy of an object's work: the number of depth samples of its rays that fall inside the unit ball at
    the INITIAL pose (what the first decoder pass of the first iteration evaluates, loss.py:60-70), computed
    analytically per ray (segment of the ray inside the ball / sample spacing).  Cheap (numpy, O(rays))."""
    T = np.asarray(detection["t_cam_obj"], np.float64).reshape(4, 4)
    rays = np.asarray(detection["rays"], np.float64).reshape(-1, 3)
    if rays.shape[0] == 0:
        return 0.0
    T_oc = np.linalg.inv(T)
    R, t = T_oc[:3, :3], T_oc[:3, 3]
    scale = np.cbrt(np.linalg.det(T[:3, :3]))
    d_min, d_max = T[2, 3] - scale, T[2, 3] + scale
    u = rays @ R.T                                    # x_o(d) = d * u + t
    a = (u * u).sum(1)
    b = 2.0 * (u @ t)
    c = float(t @ t) - 1.0
    disc = b * b - 4.0 * a * c
    ok = disc > 0
    sq = np.sqrt(np.where(ok, disc, 0.0))
    lo = np.maximum((-b - sq) / (2 * a), d_min)
    hi = np.minimum((-b + sq) / (2 * a), d_max)
    seg = np.where(ok, np.maximum(hi - lo, 0.0), 0.0)
    return float(seg.sum() / max(d_max - d_min, 1e-12) * (num_depth_samples - 1))


class RecordGatherer:
    """All-gather of per-rank record blocks with PREALLOCATED buffers (no per-step zeros / empty / cat): ``counts[r]``
    records of 82 floats from rank r, padded to the maximum.  ``gather(local)`` enqueues one
    ``all_gather_into_tensor`` on the current stream and returns the (world * max, 82) buffer; ``ordered()`` gives the
    rows in rank order without the padding (a copy only when the counts are ragged)."""

    def __init__(self, counts: Sequence[int], device, group=None):
        import torch
        import torch.distributed as dist
        self.counts = list(counts)
        self.group = group
        self.world = dist.get_world_size(group)
        self.mx = max(self.counts) if self.counts else 0
        self.pad = torch.zeros((self.mx, RECORD_LEN), dtype=torch.float32, device=device)
        self.out = torch.empty((self.world * self.mx, RECORD_LEN), dtype=torch.float32, device=device)
        self.ragged = any(c != self.mx for c in self.counts)

    def gather(self, local_records):
        import torch.distributed as dist
        src = local_records
        if local_records.shape[0] != self.mx:
            self.pad[: local_records.shape[0]].copy_(local_records)
            src = self.pad
        dist.all_gather_into_tensor(self.out, src, group=self.group)
        return self.out

    def ordered(self):
        import torch
        if not self.ragged:
            return self.out
        return torch.cat([self.out[r * self.mx: r * self.mx + self.counts[r]] for r in range(self.world)], 0)


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
