// Multi-GPU sharding of a plan, host side, no device needed: one rank per GPU, work item = (beam, contiguous
// energy-layer range). The reference is single-GPU - its beam loop (kernel_wrapper.cu:601) and layer loop (cu:916) are serial - and
// dose is linear in beams and in layers (`*res += tmp` cu:91-93, `atomicAdd` into the BEV dose kernel_wrapper.cuh:486), which is what
// makes the shards independent until the final sum.
//
// A rank that holds any layer of a beam repeats that beam's tracer, set-up and transfer in full, so the partition lays the plan's
// layers end to end and cuts them into `world` consecutive runs of near-equal cost where a run pays a fixed cost for every beam it
// touches: a beam is split only where a cut falls inside it (at most world - 1 extra beam pieces in the whole plan).
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>
#include "../../include/rtdose.h"
#include "host_geom.hpp"

namespace rtd {

struct WorkItem { int rank, beam, lo, hi; };

// Relative cost of each energy layer: rays x live depth steps x mean kernel-superposition taps. Live steps grow with the peak depth;
// the lateral sigma grows roughly linearly with depth (~2 % of range plus the in-air spot size), so
// taps ~ (2 * 3 sigma / pitch + 1)^2 averaged over depth; + 130 for the per-update cost of the depth scan, the radius passes and the
// superposition's per-(tile, layer) work that does not grow with the radius (weight table, loads, list) - refitted in round 2 to the
// per-rank stage sums of C3 on 8 B200s (profiles/r02_bench_c3_n8.json: superposition ~ 6.8e-7 ms per unit + 0.012 ms per layer).
inline std::vector<double> layerCost(const rtd_beam& b, const float* energies, const float* peakDepths, int nE) {
    const double pitch = std::min(b.raySpacing[0], b.raySpacing[1]);
    std::vector<double> out(b.nLayers);
    for (int l = 0; l < b.nLayers; ++l) {
        const double peak = vectorInterpolate(peakDepths, nE, findDecimalOrdered(energies, nE, b.energies[l]));
        const double steps = peak / std::fabs(double(b.spotDelta[2]));
        const double sigEnd = 0.022 * peak + 0.3;
        double taps = 0.0;
        for (int i = 0; i < 16; ++i) {
            const double t = double(i) / 15.0;
            const double r = std::nearbyint(3.0 * (0.3 + (sigEnd - 0.3) * t) / pitch);     // round half to even, like Python's round()
            // radii above 16 run in the two-band kernels (ks2<24> / ks2<32>: 5 / 3 warps per SM): about half the tap rate
            taps += (2.0 * r + 1.0) * (2.0 * r + 1.0) * (r > 16.0 ? 1.8 : 1.0);
        }
        out[l] = steps * (taps / 16.0 + 130.0);
    }
    return out;
}

// in layerCost units: the tracer + transfer + set-up a rank repeats for every beam it touches
// (round 2, C3 on 8 B200s: tracer + set-up + pre-convolution + depth-scan floor + radius + transfer = 0.33 ms per beam piece whatever its
// layer count, i.e. ~850 units per tracer step; round 1's kernels gave 300 - 530)
constexpr double PIECE_COST_PER_STEP = 850.0;
inline double pieceCost(const rtd_beam& b) { return PIECE_COST_PER_STEP * double(b.steps); }

// Assigns every (beam, layer) to exactly one of `world` ranks. chunksPerBeam = 0: consecutive runs of near-equal cost (see above);
// > 0: every beam is split into that many ranges, assigned with a longest-processing-time greedy (finer interleaving; forces layer splits).
inline std::vector<WorkItem> partitionPlan(const rtd_beam* beams, int nBeams, const float* energies, const float* peakDepths, int nE, int world,
                                           int chunksPerBeam = 0) {
    std::vector<WorkItem> out;
    if (world <= 1) { for (int b = 0; b < nBeams; ++b) out.push_back({0, b, 0, beams[b].nLayers}); return out; }
    std::vector<std::vector<double>> costs(nBeams);
    std::vector<double> fixed(nBeams);
    for (int b = 0; b < nBeams; ++b) { costs[b] = layerCost(beams[b], energies, peakDepths, nE); fixed[b] = pieceCost(beams[b]); }
    auto sum = [](const std::vector<double>& c, int lo, int hi) { double s = 0.0; for (int i = lo; i < hi; ++i) s += c[i]; return s; };
    if (chunksPerBeam > 0) {
        struct It { double c; int beam, lo, hi; };
        std::vector<It> items;
        for (int b = 0; b < nBeams; ++b) {
            const std::vector<double>& cost = costs[b];
            const int L = beams[b].nLayers, n = std::min(chunksPerBeam, L);
            const double target = sum(cost, 0, L) / n;
            int lo = 0, made = 0; double acc = 0.0;
            for (int l = 0; l < L; ++l) {
                acc += cost[l];
                const int remL = L - (l + 1), remC = n - made - 1;
                if ((acc >= target && remC > 0 && remL >= remC) || (remL == remC && remC > 0)) {
                    items.push_back({sum(cost, lo, l + 1), b, lo, l + 1});
                    lo = l + 1; acc = 0.0; ++made;
                }
            }
            if (lo < L) items.push_back({sum(cost, lo, L), b, lo, L});
        }
        std::stable_sort(items.begin(), items.end(), [](const It& a, const It& c) { return a.c > c.c; });
        std::vector<double> loads(world, 0.0);
        for (const It& it : items) {
            const int r = int(std::min_element(loads.begin(), loads.end()) - loads.begin());
            loads[r] += it.c;
            out.push_back({r, it.beam, it.lo, it.hi});
        }
        std::stable_sort(out.begin(), out.end(), [](const WorkItem& a, const WorkItem& c) {
            return a.rank != c.rank ? a.rank < c.rank : (a.beam != c.beam ? a.beam < c.beam : a.lo < c.lo); });
        return out;
    }
    // Greedy fill: a rank takes layers in plan order until its load (layers + one pieceCost per beam touched) reaches the target; the
    // cut goes before the layer that overshoots by more than it undershoots. The last rank takes the rest.
    auto fill = [&](double target) {
        std::vector<WorkItem> parts;
        int r = 0; double load = 0.0;
        for (int b = 0; b < nBeams; ++b) {
            const std::vector<double>& c = costs[b];
            int lo = 0; bool started = false;
            for (int l = 0; l < (int)c.size(); ++l) {
                double add = c[l] + (started ? 0.0 : fixed[b]);
                const double over = load + add - target;
                if (r < world - 1 && load > 0.0 && over > 0.0 && over > target - load) {
                    if (started) parts.push_back({r, b, lo, l});
                    ++r; load = 0.0; lo = l; started = false;
                    add = c[l] + fixed[b];
                }
                load += add;
                started = true;
            }
            parts.push_back({r, b, lo, (int)c.size()});
        }
        return parts;
    };
    auto makespan = [&](const std::vector<WorkItem>& parts) {
        std::vector<double> loads(world, 0.0);
        for (const WorkItem& it : parts) loads[it.rank] += sum(costs[it.beam], it.lo, it.hi) + fixed[it.beam];
        return *std::max_element(loads.begin(), loads.end());
    };
    double total = 0.0, maxFixed = 0.0;
    for (int b = 0; b < nBeams; ++b) { total += sum(costs[b], 0, (int)costs[b].size()) + fixed[b]; maxFixed = std::max(maxFixed, fixed[b]); }
    // the pieces created by the cuts add their own fixed cost, so the right target lies a little above total / world:
    // scan a few and keep the partition with the smallest makespan
    double bestSpan = INFINITY;
    for (int k = 0; k <= 2 * world; ++k) {
        std::vector<WorkItem> parts = fill((total + maxFixed * k * 0.5) / world);
        const double span = makespan(parts);
        if (span < bestSpan - 1e-9) { bestSpan = span; out = std::move(parts); }
    }
    return out;
}

// Conservative z-slab [z0, z1] (inclusive, dose-grid slices) outside which NO beam of the plan can deposit dose, from geometry alone,
// identical on every rank: the haloed BEV volume of each beam over ALL tracer steps (ray grid = spot grid + cutoff sigma + half a ray,
// rounded up to 32 x 8 rays, + the 32-ray halo of the superposition, kernel_wrapper.cu:612-663, kernel_wrapper.cuh:303-308) mapped
// through the divergence and gantryToDoseIdx (float3_from_fan_transform.cu:37-42). Only this slab of the ranks' partial grids is
// zeroed and summed. Returns false if the slab is empty.
inline bool doseSlab(const rtd_beam* beams, int nBeams, const int doseDims[3], float convSigmaCutoff, int& z0, int& z1) {
    const int nz = doseDims[2];
    int lo = nz, hi = -1;
    for (int bi = 0; bi < nBeams; ++bi) {
        const rtd_beam& b = beams[bi];
        double smax[2] = {0.0, 0.0};
        for (int l = 0; l < b.nLayers; ++l) { smax[0] = std::max(smax[0], double(b.sigmas[2 * l])); smax[1] = std::max(smax[1], double(b.sigmas[2 * l + 1])); }
        const double res[2] = {b.raySpacing[0], b.raySpacing[1]}, n[2] = {double(b.sx), double(b.sy)};
        const double tail[2] = {31.0, 7.0};
        double xy0[2], xy1[2];
        for (int d = 0; d < 2; ++d) {
            // the ray grid starts at ceil((first spot - cutoff sigma - res/2) / res) and is rounded UP to 32 x 8 rays at its far end;
            // 32 halo rays + 1 ray of interpolation reach + 1 of slack on both sides
            const double mLo = convSigmaCutoff * smax[d] + 0.5 * res[d] + 34.0 * res[d];
            const double mHi = convSigmaCutoff * smax[d] + 0.5 * res[d] + (tail[d] + 34.0) * res[d];
            const double a = b.spotOffset[d], e = b.spotOffset[d] + (n[d] - 1.0) * b.spotDelta[d];
            xy0[d] = std::min(a, e) - mLo; xy1[d] = std::max(a, e) + mHi;
        }
        const double dz = b.spotDelta[2], zs = b.spotOffset[2];
        const double gzs[2] = {zs - dz, zs + (double(b.steps) + 1.0) * dz};
        for (double gz : gzs) {
            const double sc[2] = {1.0 - gz / double(b.sourceDist[0]), 1.0 - gz / double(b.sourceDist[1])};     // inf -> 1 (parallel beam)
            for (int cx = 0; cx < 2; ++cx)
                for (int cy = 0; cy < 2; ++cy) {
                    const double gx = (cx ? xy1[0] : xy0[0]) * sc[0], gy = (cy ? xy1[1] : xy0[1]) * sc[1];
                    const double pz = double(b.gantryToDoseIdxM[6]) * gx + double(b.gantryToDoseIdxM[7]) * gy + double(b.gantryToDoseIdxM[8]) * gz + double(b.gantryToDoseIdxV[2]);
                    lo = std::min(lo, int(std::floor(pz)) - 2); hi = std::max(hi, int(std::ceil(pz)) + 2);
                }
        }
    }
    lo = std::max(lo, 0); hi = std::min(hi, nz - 1);
    z0 = lo; z1 = hi;
    return hi >= lo;
}

}  // namespace rtd
