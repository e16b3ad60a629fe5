"""Host-side plan of one spectral step: batch sharding, the packed all-reduce payload and the normalisers.

Pure host logic (no CUDA, no kernels) so that the N > 1 path can be exercised on CPU with the ``gloo``
backend (``tests/test_dist_cpu.py``).  The reference computes its loss means and its PCA over the whole
batch on one device (``train_pkl_file_task_aware.py:597-629``, ``pkl_file_model_updated.py:489-511``); the
sharded form reproduces that exactly (up to summation order) because everything exchanged is a plain sum:

    packed = [ 2S loss sums | per view: n, sum_b z (d), sum_b z z^T (d*d) ]      float64, ONE all-reduce (sum)

and the global normalisers ``1 / (S * count_s * world)`` are known from shapes alone, so neither the loss
value nor the input gradient needs a second collective.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence, Tuple

import torch


def num_frames(T: int, n_fft: int, hop: int, center: bool) -> int:
    Tp = T + (2 * (n_fft // 2) if center else 0)
    return 0 if Tp < n_fft else 1 + (Tp - n_fft) // hop


def shard_bounds(batch: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous batch shard of ``rank``: items [lo, hi).  All views of an item stay on one rank (rows of the
    audio tensor are item-major, view-minor), so correlated views are never split."""
    if not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world of {world}")
    base, rem = divmod(batch, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


@dataclass
class StepPlan:
    rows: int            # clips on THIS rank (items * views * channels)
    T: int
    n_ffts: Sequence[int]
    hops: Sequence[int]
    n_views: int
    d: int
    world: int = 1
    global_rows: int = 0  # clips on ALL ranks; 0 = equal shards (rows * world).  Uneven shards (shard_bounds) pass the true sum

    def __post_init__(self):
        self.n_ffts, self.hops = list(self.n_ffts), list(self.hops)
        assert len(self.n_ffts) == len(self.hops)                       # msstftd.py:177
        self.S = len(self.n_ffts)
        self.stat_stride = 1 + self.d + self.d * self.d
        self.counts = [self.rows * (n // 2 + 1) * num_frames(self.T, n, h, False) for n, h in zip(self.n_ffts, self.hops)]
        if min(self.counts) <= 0:
            raise ValueError(f"signal of {self.T} samples is shorter than the largest n_fft")
        if self.global_rows <= 0:
            self.global_rows = self.rows * self.world
        if self.global_rows < self.rows:
            raise ValueError(f"global_rows {self.global_rows} < local rows {self.rows}")
        self.global_counts = [self.global_rows * (n // 2 + 1) * num_frames(self.T, n, h, False)
                              for n, h in zip(self.n_ffts, self.hops)]

    # ---- packed buffer layout
    @property
    def packed_len(self) -> int:
        return 2 * self.S + self.n_views * self.stat_stride

    def sums_slice(self) -> slice:
        return slice(0, 2 * self.S)

    def stats_slice(self) -> slice:
        return slice(2 * self.S, self.packed_len)

    def views(self, packed: torch.Tensor):
        """(sums [2S], stats [nv, 1 + d + d*d]) as views of the packed buffer."""
        if packed.numel() != self.packed_len or packed.dtype != torch.float64:
            raise ValueError(f"packed buffer must be float64 [{self.packed_len}]")
        return packed[self.sums_slice()], packed[self.stats_slice()].view(self.n_views, self.stat_stride)

    # ---- normalisers (GLOBAL element counts: rows of all ranks, equal shards or not)
    def inv_counts(self) -> List[float]:
        """d loss / d sums: loss = sum_j sums[j] * inv[j] with the GLOBAL element counts."""
        return [1.0 / (self.S * c) for c in self.global_counts for _ in (0, 1)]

    def loss_from_sums(self, sums: torch.Tensor) -> torch.Tensor:
        return torch.dot(sums.double().cpu(), torch.tensor(self.inv_counts(), dtype=torch.float64))

    # ---- the exchange
    def allreduce(self, packed: torch.Tensor, group=None, what: str = "all") -> None:
        """In-place sum over ranks of the packed buffer (or one of its two slices, so that the eigen-solve, which
        needs the statistics only, can start while the loss kernel still runs)."""
        if group is None and self.world == 1:
            return
        import torch.distributed as dist
        sl = {"all": slice(0, self.packed_len), "sums": self.sums_slice(), "stats": self.stats_slice()}[what]
        dist.all_reduce(packed[sl], op=dist.ReduceOp.SUM, group=group)
