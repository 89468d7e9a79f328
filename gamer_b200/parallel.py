"""Multi-GPU partitioning of the render path: one process per GPU (torch.distributed), no collective in
the data path except the final gather of disjoint tiles.

The reference is single-process; its two scheduling strategies are the shuffled render list
(rasterizer.cpp:115-122) and RasterThread's equal chunks of that list (rasterthread.cpp:9-24).
The multi-GPU analogue here:
  * frames / galaxies: unit k -> rank k mod world (no exchange at all);
  * image tiles: interleaved bands of BAND rows (band b -> rank b mod world), which balances the very
    uneven cost per row of edge-on views the way the reference's shuffle does, followed by ONE gather
    of the disjoint bands to rank 0 (NCCL over NVLink on GPUs; gloo in the CPU tests).
Everything in this module is index arithmetic and torch plumbing and runs on CPU tensors as well.
"""
from typing import List, Tuple

import torch
import torch.distributed as dist

BAND = 32
_row_index_cache = {}


def _row_index(r, world, height, band, device):
    key = (r, world, height, band, str(device))
    if key not in _row_index_cache:
        _row_index_cache[key] = torch.as_tensor(band_rows(r, world, height, band), device=device, dtype=torch.long)
    return _row_index_cache[key]


def units_of(rank: int, world: int, n_units: int) -> List[int]:
    """frame / galaxy sharding: unit k belongs to rank k mod world."""
    return list(range(rank, n_units, world))


def band_rows(rank: int, world: int, height: int, band: int = BAND) -> List[int]:
    """image rows owned by `rank`, in the order its local framebuffer stores them."""
    rows = []
    for b in range(rank, (height + band - 1) // band, world):
        rows.extend(range(b * band, min(height, (b + 1) * band)))
    return rows


def band_tile(scene, rank: int, world: int, band: int = BAND):
    """The rank's share of `scene` (a gamer_b200.model.Scene covering the full frame) as a tile scene
    using GamerFrame's band_rows/band_stride addressing."""
    f = scene.frame
    n = len(band_rows(rank, world, f.height, band))
    if world == 1:
        return scene.tile(f.x0, f.y0, f.width, f.height)
    if n == 0:          # more ranks than bands (a small frame): this rank owns nothing and neither uploads nor renders
        return None
    return scene.tile(f.x0, f.y0 + rank * band, f.width, n, band_rows=band, band_stride=band * world)


def gather_bands(local: torch.Tensor, height: int, rank: int, world: int, band: int = BAND, dst: int = 0):
    """Gather every rank's band-ordered rows [rows_r, ...] to `dst` and return the frame [height, ...]
    in image order there (None elsewhere). Ranks may own different row counts (ragged last band)."""
    if world == 1:
        return local
    counts = [len(band_rows(r, world, height, band)) for r in range(world)]
    pad = max(counts)
    if local is None:                               # a rank without rows (band_tile returned None) still takes part
        raise ValueError("gather_bands needs a tensor on every rank: pass an empty one (shape [0, ...]) for a rank without rows")
    if local.shape[0] != pad:                       # gather needs equal shapes: pad the short ranks
        padded = local.new_zeros((pad,) + tuple(local.shape[1:]))
        padded[: local.shape[0]] = local
        local = padded
    bufs = [torch.empty_like(local) for _ in range(world)] if rank == dst else None
    dist.gather(local, bufs, dst=dst)
    if rank != dst:
        return None
    out = local.new_empty((height,) + tuple(local.shape[1:]))
    for r in range(world):
        out.index_copy_(0, _row_index(r, world, height, band, local.device), bufs[r][: counts[r]])
    return out


def job_cost(scene) -> float:
    """Cost proxy of one render job that needs no rendering to evaluate: pixels x the noise octaves its
    active components would run at an in-slab sample (10 disk / stars, 9 dust / dust2 / dust positive; a bulge counts 1)."""
    octaves = {0: 1, 1: 10, 2: 9, 3: 9, 4: 9, 5: 10, 6: 0}
    per_sample = sum(octaves[int(c.class_id)] for c in scene.components if c.active == 1.0)
    return float(scene.frame.width) * float(scene.frame.height) * per_sample


def longest_job_first(costs: List[float], world: int) -> List[List[int]]:
    """Greedy longest-processing-time assignment of independent jobs (galaxies of a batch, frames of a sequence) to
    ranks: jobs in descending cost, each to the least loaded rank so far. Deterministic (ties by job index), so every
    rank computes the same table without communication. The reference's queue runs its jobs one at a time in
    submission order (renderqueue.cpp:63-87); with several GPUs the order is what balances them.
    Returns the job indices of every rank, in execution order."""
    order = ljf_order(costs)
    load = [0.0] * world
    jobs = [[] for _ in range(world)]
    for j in order:
        r = min(range(world), key=lambda q: (load[q], q))
        jobs[r].append(j)
        load[r] += costs[j]
    return jobs


class JobQueue:
    """One queue of independent jobs for all ranks, the multi-GPU form of the reference's RenderQueue, which hands its
    jobs to whichever renderer is free (renderqueue.cpp:63-87). `order` is the execution order every rank computes
    for itself (longest job first: longest_job_first's sort); iterating yields the next index of it that no rank has
    taken yet, each exactly once over the whole job. The hand-out is one atomic add on the process group's store (the
    rendezvous TCPStore: ~0.1 ms, against milliseconds per job) - control plane only, no pixel ever goes through it.
    With one rank, or no process group, it is a local counter. A static table (longest_job_first) balances the ESTIMATED
    costs; the queue balances the real ones: whichever rank finishes early takes the next job."""

    def __init__(self, order: List[int], name: str):
        self.order, self.key, self.local = list(order), "gamer_b200/jobqueue/" + name, 0
        self.store = None
        if dist.is_initialized() and dist.get_world_size() > 1:
            self.store = dist.distributed_c10d._get_default_store()

    def __iter__(self):
        return self

    def __next__(self) -> int:
        if self.store is not None:
            i = self.store.add(self.key, 1) - 1
        else:
            i, self.local = self.local, self.local + 1
        if i >= len(self.order):
            raise StopIteration
        return self.order[i]


def ljf_order(costs: List[float]) -> List[int]:
    """Job indices in descending cost, ties by index: the order longest_job_first assigns in and JobQueue hands out in."""
    return sorted(range(len(costs)), key=lambda j: (-costs[j], j))


def max_over_ranks(values: List[float], device) -> List[float]:
    t = torch.tensor(values, dtype=torch.float64, device=device)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.tolist()


def sum_over_ranks(values: List[float], device) -> List[float]:
    t = torch.tensor(values, dtype=torch.float64, device=device)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.tolist()
