"""Multi-GPU sharding of a plan: one process per GPU, work item = (beam, contiguous energy-layer
range), balanced with a longest-processing-time greedy; every rank accumulates its items into a local full-size
partial dose grid and the grids are summed with one NCCL reduce over NVLink/NVSwitch (``torch.distributed``).

The reference is single-GPU (its beam loop, kernel_wrapper.cu:601, and layer loop, cu:916, are serial); dose is
linear in beams and in layers (``*res += tmp`` cu:91-93, ``atomicAdd`` into the BEV dose kernel_wrapper.cuh:486),
which is what makes the shards independent until the final sum.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np

from .cases import Beam, find_decimal_ordered, vector_interpolate

WorkItem = Tuple[int, int, int]   # (beam number, first layer, one-past-last layer)


def layer_cost(beam: Beam, luts) -> np.ndarray:
    """Relative cost of each energy layer: rays x live depth steps x mean kernel-superposition taps.
    Live steps grow with the peak depth; the lateral sigma grows roughly linearly with depth (~2 % of range plus
    the in-air spot size), so taps ~ (2*3*sigma/pitch + 1)^2 averaged over depth."""
    pitch = float(min(beam.ray_spacing))
    out = np.empty(beam.n_layers, np.float64)
    for l in range(beam.n_layers):
        peak = float(vector_interpolate(luts.peakDepths, find_decimal_ordered(luts.energiesPerU, beam.energies[l])))
        steps = peak / abs(float(beam.spot_delta[2]))
        sig_end = 0.022 * peak + 0.3
        taps = np.mean([(2.0 * round(3.0 * (0.3 + (sig_end - 0.3) * t) / pitch) + 1.0) ** 2 for t in np.linspace(0.0, 1.0, 16)])
        out[l] = steps * (taps + 40.0)   # + per-update cost of the IDD/sigma and radius passes
    return out


def partition(beams: Sequence[Beam], luts, world_size: int, chunks_per_beam: int = 0) -> List[List[WorkItem]]:
    """Split every beam into contiguous layer ranges of near-equal cost and assign them to ``world_size`` ranks
    (LPT greedy). Returns one list of work items per rank; every (beam, layer) appears exactly once."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    if world_size == 1:
        return [[(b, 0, beams[b].n_layers) for b in range(len(beams))]]
    nb = len(beams)
    if chunks_per_beam <= 0:
        # enough items for balance, few enough that the per-item tracer recomputation stays small
        chunks_per_beam = max(1, -(-2 * world_size // max(nb, 1)))
    items = []
    for b, beam in enumerate(beams):
        cost = layer_cost(beam, luts)
        n = min(chunks_per_beam, beam.n_layers)
        target = cost.sum() / n
        lo, acc, made = 0, 0.0, 0
        for l in range(beam.n_layers):
            acc += cost[l]
            remaining_layers = beam.n_layers - (l + 1)
            remaining_chunks = n - made - 1
            if (acc >= target and remaining_chunks > 0 and remaining_layers >= remaining_chunks) or remaining_layers == remaining_chunks and remaining_chunks > 0:
                items.append((cost[lo:l + 1].sum(), (b, lo, l + 1)))
                lo, acc, made = l + 1, 0.0, made + 1
        if lo < beam.n_layers:
            items.append((cost[lo:].sum(), (b, lo, beam.n_layers)))
    items.sort(key=lambda t: -t[0])
    loads = [0.0] * world_size
    out: List[List[WorkItem]] = [[] for _ in range(world_size)]
    for c, it in items:
        r = int(np.argmin(loads))
        loads[r] += c
        out[r].append(it)
    for r in range(world_size):
        out[r].sort()
    return out


def shard_beams(beams: Sequence[Beam], items: Sequence[WorkItem]) -> List[Beam]:
    """The ``Beam`` objects (with layer_begin/layer_end set) a rank hands to ``rtd_dose``."""
    return [beams[b].shard(lo, hi) if (lo, hi) != (0, beams[b].n_layers) else beams[b] for b, lo, hi in items]


def reduce_dose(dose_tensor, dst: int = 0, all_ranks: bool = False):
    """Sum the ranks' partial dose grids in place (``torch.distributed``; NCCL for CUDA tensors, gloo for CPU ones).
    ``all_ranks``: all-reduce instead of a reduce to ``dst``."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return dose_tensor
    if all_ranks:
        dist.all_reduce(dose_tensor, op=dist.ReduceOp.SUM)
    else:
        dist.reduce(dose_tensor, dst=dst, op=dist.ReduceOp.SUM)
    return dose_tensor
