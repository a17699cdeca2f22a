"""Multi-GPU sharding of a plan: one process per GPU, work item = (beam, contiguous energy-layer
range). The partitioner, the reachable dose slab, the stream barriers and the sum of the partial doses live in the library
(``rtd_partition``, ``rtd_dose_slab``, ``rtd_group_*``, ``rtd_plan_dose_group``); this module is the Python view of the first two
plus the NCCL alternative for the sum (``reduce_dose``, ``torch.distributed``).

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
    """The cost model of ``rtd_partition`` (csrc/sharding.hpp ``layerCost``), restated: relative cost of each energy layer =
    live depth steps x (mean kernel-superposition taps + 40). Live steps grow with the peak depth; the lateral sigma grows roughly
    linearly with depth (~2 % of range plus the in-air spot size), so taps ~ (2*3*sigma/pitch + 1)^2 averaged over depth."""
    pitch = float(min(beam.ray_spacing))
    out = np.empty(beam.n_layers, np.float64)
    for l in range(beam.n_layers):
        peak = float(vector_interpolate(luts.peakDepths, find_decimal_ordered(luts.energiesPerU, beam.energies[l])))
        steps = peak / abs(float(beam.spot_delta[2]))
        sig_end = 0.022 * peak + 0.3
        radii = [round(3.0 * (0.3 + (sig_end - 0.3) * t) / pitch) for t in np.linspace(0.0, 1.0, 16)]
        taps = np.mean([(2.0 * r + 1.0) ** 2 * (1.8 if r > 16 else 1.0) for r in radii])     # two-band kernels above radius 16: half the tap rate
        out[l] = steps * (taps + 130.0)  # + per-update cost of the IDD/sigma, radius passes and the radius-independent part of the superposition
    return out


PIECE_COST_PER_STEP = 850.0   # in layer_cost units: the tracer + transfer + setup a rank repeats for every beam it touches
                              # (round 2 fit to the per-rank stage sums of C3 on 8 B200s: 0.33 ms per beam piece)


def piece_cost(beam: Beam) -> float:
    """Cost, in ``layer_cost`` units, of the per-beam stages every rank holding ANY layer of the beam runs in full
    (fan-beam tracer over all steps, spot pre-convolution set-up, BEV -> dose-grid transfer); ``pieceCost`` in csrc/sharding.hpp."""
    return PIECE_COST_PER_STEP * float(beam.steps)


def partition(beams: Sequence[Beam], luts, world_size: int, chunks_per_beam: int = 0) -> List[List[WorkItem]]:
    """Assign every (beam, layer) to exactly one of ``world_size`` ranks as (beam, contiguous layer range) work items.

    The partitioner itself is the library's (``rtd_partition``, csrc/sharding.hpp - the one ``rtd_plan_dose_group`` uses, so that a
    C++ caller of include/rtdose.h gets the same shards); this wrapper only regroups its items per rank.
    Default (``chunks_per_beam`` = 0): the beams' layers are laid end to end and cut into ``world_size`` consecutive runs of
    near-equal cost, where a run pays a fixed cost for every beam it touches (a rank that holds any layer of a beam repeats that
    beam's tracer, set-up and transfer in full): a beam is split only where a cut falls inside it.
    ``chunks_per_beam`` > 0: every beam is split into that many ranges, assigned with an LPT greedy (forces layer splits)."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    from . import engine
    out: List[List[WorkItem]] = [[] for _ in range(world_size)]
    for rank, b, lo, hi in engine.partition(beams, luts, world_size, chunks_per_beam):
        out[rank].append((b, lo, hi))
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


def dose_slab(beams: Sequence[Beam], dose_dims, sigma_cutoff: float = 3.0) -> Tuple[int, int]:
    """Conservative z-slab [z0, z1] (inclusive, dose-grid slices) outside which NO beam of the plan can deposit dose,
    from geometry alone, identical on every rank (``rtd_dose_slab``, csrc/sharding.hpp): the haloed BEV volume of each beam over
    ALL tracer steps mapped through the divergence and gantryToDoseIdx. Only this slab of the partial dose grids is zeroed and
    summed. ``sigma_cutoff`` is rtd_opts.convSigmaCutoff (CONV_SIGMA_CUTOFF, 3 by default): it sizes the ray grid."""
    from . import engine
    return engine.dose_slab(beams, dose_dims, sigma_cutoff)
