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


PIECE_COST_PER_STEP = 400.0   # in layer_cost units: the tracer + transfer + setup a rank repeats for every beam it touches
                              # (least-squares fit of the per-rank stage totals of C3 and C4 at N = 2 and 8 on B200: 300 - 530)


def piece_cost(beam: Beam) -> float:
    """Cost, in ``layer_cost`` units, of the per-beam stages every rank holding ANY layer of the beam runs in full
    (fan-beam tracer over all steps, spot pre-convolution set-up, BEV -> dose-grid transfer)."""
    return PIECE_COST_PER_STEP * float(beam.steps)


def partition(beams: Sequence[Beam], luts, world_size: int, chunks_per_beam: int = 0) -> List[List[WorkItem]]:
    """Assign every (beam, layer) to exactly one of ``world_size`` ranks as (beam, contiguous layer range) work items.

    Default (``chunks_per_beam`` = 0): the beams' layers are laid end to end and cut into ``world_size`` consecutive
    runs of near-equal cost (McNaughton's wrap-around rule), where a run pays ``piece_cost`` for every beam it
    touches. A beam is split only where a cut falls inside it, so a rank repeats the tracer / transfer of as few beams
    as possible (at most world_size - 1 extra beam pieces in the whole plan).
    ``chunks_per_beam`` > 0: every beam is split into that many ranges, assigned with an LPT greedy (finer interleaving;
    used by the tests to force layer splits)."""
    if world_size < 1:
        raise ValueError("world_size must be >= 1")
    if world_size == 1:
        return [[(b, 0, beams[b].n_layers) for b in range(len(beams))]]
    if chunks_per_beam <= 0:
        return _partition_contiguous(beams, luts, world_size)
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


def _partition_contiguous(beams: Sequence[Beam], luts, world_size: int) -> List[List[WorkItem]]:
    costs = [layer_cost(b, luts) for b in beams]
    fixed = [piece_cost(b) for b in beams]

    def fill(target: float) -> List[List[WorkItem]]:
        """Greedy: a rank takes layers in plan order until its load (layers + one piece_cost per beam touched) reaches
        ``target``; the cut goes before the layer that overshoots by more than it undershoots. The last rank takes the
        rest. Ranks left with nothing stay empty."""
        out: List[List[WorkItem]] = [[] for _ in range(world_size)]
        r, load = 0, 0.0
        for b, c in enumerate(costs):
            lo = 0
            started = False
            for l in range(len(c)):
                add = c[l] + (0.0 if started else fixed[b])
                over = load + add - target
                if r < world_size - 1 and load > 0.0 and over > 0.0 and over > target - load:
                    if started:
                        out[r].append((b, lo, l))
                    r, load, lo, started = r + 1, 0.0, l, False
                    add = c[l] + fixed[b]
                load += add
                started = True
            out[r].append((b, lo, len(c)))
        return out

    def makespan(parts) -> float:
        return max(sum(costs[b][lo:hi].sum() + fixed[b] for b, lo, hi in items) for items in parts)

    total = sum(float(c.sum()) for c in costs) + sum(fixed)
    best, best_span = None, np.inf
    # the pieces created by the cuts add their own fixed cost, so the right target lies a little above total / world:
    # scan a few and keep the partition with the smallest makespan
    for k in range(0, 2 * world_size + 1):
        extra = sum(sorted(fixed, reverse=True)[:1]) * k * 0.5
        parts = fill((total + extra) / world_size)
        span = makespan(parts)
        if span < best_span - 1e-9:
            best, best_span = parts, span
    return best


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
    from geometry alone, identical on every rank: the haloed BEV volume of each beam over ALL tracer steps (ray grid =
    spot grid + 3 sigma + half a ray, rounded up to 32 x 8 rays, + the 32-ray halo of the superposition,
    kernel_wrapper.cu:612-663, kernel_wrapper.cuh:303-308) mapped through the divergence and gantryToDoseIdx
    (float3_from_fan_transform.cu:37-42). Only this slab of the partial dose grids is zeroed and reduced over NCCL.
    ``sigma_cutoff`` is rtd_opts.convSigmaCutoff (CONV_SIGMA_CUTOFF, 3 by default): it sizes the ray grid."""
    nx, ny, nz = dose_dims
    lo, hi = nz, -1
    for b in beams:
        L, sy, sx = b.weights.shape
        res = np.array(b.ray_spacing, np.float64)
        smax = np.max(np.asarray(b.sigmas, np.float64), axis=0)
        n = np.array([sx, sy], np.float64)
        d = np.array(b.spot_delta[:2], np.float64); o = np.array(b.spot_offset[:2], np.float64)
        # the ray grid starts at ceil((first spot - cutoff sigma - res/2) / res) and is rounded UP to 32 x 8 rays at its far end;
        # 32 halo rays + 1 ray of interpolation reach + 1 of slack on both sides
        m_lo = sigma_cutoff * smax + 0.5 * res + 34.0 * res
        m_hi = sigma_cutoff * smax + 0.5 * res + (np.array([31.0, 7.0]) + 34.0) * res
        xy0 = np.minimum(o, o + (n - 1) * d) - m_lo
        xy1 = np.maximum(o, o + (n - 1) * d) + m_hi
        dz = float(b.spot_delta[2]); z0 = float(b.spot_offset[2])
        m = np.asarray(b.gantry_to_dose_m, np.float64); v = np.asarray(b.gantry_to_dose_v, np.float64)
        sd = np.array(b.source_dist, np.float64)
        for gz in (z0 - dz, z0 + (b.steps + 1) * dz):
            scale = 1.0 - gz / sd                                       # inf -> 1 (parallel beam)
            for gx in (xy0[0], xy1[0]):
                for gy in (xy0[1], xy1[1]):
                    p = m @ np.array([gx * scale[0], gy * scale[1], gz]) + v
                    lo = min(lo, int(np.floor(p[2])) - 2); hi = max(hi, int(np.ceil(p[2])) + 2)
    lo, hi = max(lo, 0), min(hi, nz - 1)
    if hi < lo:
        return 0, -1
    return lo, hi
