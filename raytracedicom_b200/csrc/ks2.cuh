// K3b (second generation): kernel superposition, layer-fused and warp-autonomous.
//
// Reference: kernelSuperposition<rad> (/root/reference/src/kernel_wrapper.cuh:432-489) is launched per layer and per
// radius batch; every tile does (2r+1)^2 volatile shared-memory read-modify-writes per voxel with a block barrier per
// row offset and then (32+2r)(8+2r) global float atomics.  Here:
//   * work item = (depth slice kk, 32x8 tile); ONE WARP owns the item and loops over ALL energy layers of the beam,
//     accumulating them into a warp-private shared-memory patch, so the haloed BEV dose receives one global reduction
//     per item instead of one per (item, layer): ~L times fewer global atomics, and no block barriers at all;
//   * lane x owns the source column x of the tile.  The erf-difference weights of its 8 voxels (same recurrence as
//     cuh:459-467; erf itself is ks_erf_pos below, a single-range fit with 1e-7 absolute error instead of libdevice's erff -
//     a deliberate, bounded numerics deviation) go through a warp-private table so that the hot loop is a SMALL rolled loop
//     over the column offset j (instruction-cache resident) with the row weights e[y][|i|] held in registers:
//     per j the lane accumulates a whole output column (P+2RB rows) in registers - P(2RB+1) FMAs as packed FFMA2 over row
//     pairs, no memory traffic -
//     and adds it to the patch columns x+j and x-j (the horizontal weight is symmetric: +j and -j share the FMAs); the
//     patch is column-major so that this read-modify-write is 128-bit wide (see ks2_col_pitch);
//   * the radius applied to a tile is still the reference's BATCH radius (cuh:443-448 vs 459-473).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "kernels.cuh"

// Measurement switches of three round-2 micro-optimisations; tools/gpu_r2_c.sh builds the variants (profiles/r02_ks2_variants_ab.json,
// C3 superposition per plan): baseline 1.123 ms; weight table with packed fp32 1.093 ms (kept); shared-memory words of the next
// column offset requested right after the FMA block 1.120 ms (neutral, off); next work item requested one item ahead 1.411 ms
// (25 % SLOWER, off - the kernel is bound by shared-memory bandwidth and by latency at 2 warps per scheduler, not by the atomic).
#ifndef RTD_KS2_PACKED_ERF
#define RTD_KS2_PACKED_ERF 1
#endif
#ifndef RTD_KS2_EARLY_G
#define RTD_KS2_EARLY_G 0
#endif
#ifndef RTD_KS2_R0_FAST
#define RTD_KS2_R0_FAST 0      // measured (tools/gpu_r2_l.sh): C3 superposition 1.096 -> 1.104 ms, C4 12.55 -> 12.60: no gain, off
#endif
#ifndef RTD_KS2_ITEM_PREFETCH
#define RTD_KS2_ITEM_PREFETCH 0
#endif

namespace rtd {

struct Ks2Params {
    const float* idd; const float* rsig; float* bev;
    const int* batchRad;            // [L][34] radius applied to a tile of own radius r
    const uint8_t* radMap;          // [L][nz][tiles] own radius, 255 = not live
    const int* layerFirstPassive;   // [L]
    const BeamDev* bd;              // slices per layer (nz) and first slice (bfi): known to the device only (beam_setup_kernel)
    int L, Rx, Ry, tilesX, tiles;
    const int* bigCount; const uint32_t* bigItems;   // items with a layer of applied radius > 16 (big_items_kernel)
    int bodyStep;                   // 1, 2 or 4: radii <= 16 run in the unrolled body of the next multiple (see ks2_kernel);
                                    // 0 = 2 if the beam has items with a radius > 16 (*bigCount > 0), else 1
};

// erf(x) for x >= 0 as 1 - 2^(x Q(x)), Q = degree-7 fit of log2(erfc(x))/x on [0,4] (erf(4) rounds to 1 in fp32).
// Max abs error 1.0e-7 (+ ex2.approx's 2 ulp), i.e. the fp32 resolution of erf near 1; 11 instructions against
// ~25 for libdevice's two-range erff, which dominated the instruction count of this kernel at small radii.
// The weights only need erf differences to ~1e-6 of the central weight (BASELINE.json bound: 1e-3 of max dose).
__device__ __forceinline__ float ks_erf_pos(float x) {
    x = fminf(x, 4.0f);
    float q = -4.223215728e-05f;
    q = fmaf(q, x, 4.191363114e-04f);
    q = fmaf(q, x, -1.401739311e-03f);
    q = fmaf(q, x, -9.219333297e-04f);
    q = fmaf(q, x, 2.838649414e-02f);
    q = fmaf(q, x, -1.485444903e-01f);
    q = fmaf(q, x, -9.184023738e-01f);
    q = fmaf(q, x, -1.627909660e+00f);
    return 1.0f - exp2f(x * q);
}

// The weight table costs as many instructions as the taps for applied radii <= 6: two voxels (rows y, y + 1 of a lane) are evaluated
// per instruction with packed fp32. h(x) = erf(x) / 2 = 0.5 - 0.5 * 2^(x Q(x)); min and ex2 have no packed form.
typedef unsigned long long ks_u64;
__device__ __forceinline__ ks_u64 ks_pk(float lo, float hi) { ks_u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void ks_upk(ks_u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ ks_u64 ks_fma2(ks_u64 a, ks_u64 b, ks_u64 c) { ks_u64 d; asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ ks_u64 ks_add2(ks_u64 a, ks_u64 b) { ks_u64 d; asm("add.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ ks_u64 ks_mul2(ks_u64 a, ks_u64 b) { ks_u64 d; asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ ks_u64 ks_bc(float v) { return ks_pk(v, v); }
__device__ __forceinline__ ks_u64 ks_half_erf_pos2(ks_u64 x2) {
    float x0, x1; ks_upk(x2, x0, x1);
    const ks_u64 x = ks_pk(fminf(x0, 4.0f), fminf(x1, 4.0f));
    ks_u64 q = ks_bc(-4.223215728e-05f);
    q = ks_fma2(q, x, ks_bc(4.191363114e-04f));
    q = ks_fma2(q, x, ks_bc(-1.401739311e-03f));
    q = ks_fma2(q, x, ks_bc(-9.219333297e-04f));
    q = ks_fma2(q, x, ks_bc(2.838649414e-02f));
    q = ks_fma2(q, x, ks_bc(-1.485444903e-01f));
    q = ks_fma2(q, x, ks_bc(-9.184023738e-01f));
    q = ks_fma2(q, x, ks_bc(-1.627909660e+00f));
    float a0, a1; ks_upk(ks_mul2(x, q), a0, a1);
    return ks_fma2(ks_pk(exp2f(a0), exp2f(a1)), ks_bc(-0.5f), ks_bc(0.5f));
}

constexpr int KS2_MAXRB = 32;       // largest batch radius (RTD_MAX_SUPERP_R)
constexpr int KS2_MAXL = 256;

// Warp-private patch, COLUMN-major: patch[X][Y], X = tile column + RBMAX (32 + 2 RBMAX columns), Y = tile row + RBMAX
// (8 + 2 RBMAX rows, a multiple of 4), column pitch padded so that 128-bit accesses of 8 consecutive lanes (one column
// each) fall into 8 distinct groups of 4 banks: pitch % 4 == 0 and pitch % 32 in {4, 12, 20, 28}  ->  44 (RBMAX 16), 28 (8).
// A lane's register-accumulated output column is then added with LDS.128 / STS.128: a quarter of the shared-memory
// instructions of a row-major patch at the same bank-conflict-free bandwidth.
// The 4 pad floats at the end of every column hold the item's layer list (8 uint16 entries per column).
__host__ __device__ constexpr int ks2_rows(int rbmax) { return 8 + 2 * rbmax; }
__host__ __device__ constexpr int ks2_cols(int rbmax) { return 32 + 2 * rbmax; }
__host__ __device__ constexpr int ks2_col_pitch(int rbmax) {
    int p = ks2_rows(rbmax) + 4;
    while (p % 4 != 0 || (p % 32) % 8 != 4) ++p;
    return p;
}
// per-warp shared memory (floats): patch (with the layer list in its pad) + weight table
__host__ __device__ constexpr int ks2_patch_floats(int rbmax) { return ks2_cols(rbmax) * ks2_col_pitch(rbmax); }
__host__ __device__ constexpr int ks2_table_floats(int rbmax) { return 8 * (rbmax + 1) * 32; }
__host__ __device__ constexpr int ks2_warp_floats(int rbmax) { return ks2_patch_floats(rbmax) + ks2_table_floats(rbmax); }
static_assert(ks2_col_pitch(32) == 76 && ks2_col_pitch(24) == 60 && ks2_col_pitch(16) == 44 && ks2_col_pitch(8) == 28, "column pitch");
static_assert((ks2_col_pitch(16) - ks2_rows(16)) * 2 * ks2_cols(16) >= KS2_MAXL && (ks2_col_pitch(8) - ks2_rows(8)) * 2 * ks2_cols(8) >= KS2_MAXL &&
              (ks2_col_pitch(32) - ks2_rows(32)) * 2 * ks2_cols(32) >= KS2_MAXL && (ks2_col_pitch(24) - ks2_rows(24)) * 2 * ks2_cols(24) >= KS2_MAXL,
              "layer list fits the column pads");

// Packed fp32 (sm_100: fma.rn.f32x2 -> FFMA2, add.f32x2 -> FADD2; two IEEE fp32 operations per issue slot). The superposition is
// bound by instruction issue and shared-memory wavefronts, not by the FMA pipe (profiles/r01_ncu_full_ks2.txt: issue-active 56 %,
// FMA pipe 35 %), so halving the issue slots of the FMA block and of the patch read-modify-write is what pays.
__device__ __forceinline__ void ks_fma_lo(ks_u64& c, float a, float b) { float lo, hi; ks_upk(c, lo, hi); lo = fmaf(a, b, lo); c = ks_pk(lo, hi); }
__device__ __forceinline__ void ks_fma_hi(ks_u64& c, float a, float b) { float lo, hi; ks_upk(c, lo, hi); hi = fmaf(a, b, hi); c = ks_pk(lo, hi); }

// The 8 source voxels of a lane (rows y = 0..7 of its column) with one vertical weight each, packed by ROW PAIRS: the voxels of rows
// y and y + 1 at the same vertical offset land in adjacent output rows, so {out[Y], out[Y+1]} += {g[y], g[y+1]} * {e[y], e[y+1]} is one
// FFMA2 with no duplicated operand. The output column is held as 64-bit pairs aligned with the 128-bit words of the patch column
// ("window" index w, pair m = w / 2); a vertical offset whose row-0 output has an even window index pairs rows (0,1)(2,3)(4,5)(6,7),
// an odd one pairs (1,2)(3,4)(5,6) and leaves rows 0 and 7 to scalar FFMAs on the halves of their pairs. The parity is a compile-time
// property of the weight index, so each weight is stored once, in the pairing its parity needs: 8 registers per weight index as before.
struct Ks2W {              // dose x vertical weight of one vertical offset for the lane's 8 voxels
    ks_u64 p[4];           // even parity: rows (0,1)(2,3)(4,5)(6,7); odd parity: rows (1,2)(3,4)(5,6), p[3] unused
    float s0, s7;          // odd parity: rows 0 and 7
};
struct Ks2G {              // horizontal weight of the lane's 8 voxels for one column offset, in both pairings
    ks_u64 e[4], o[3]; float s0, s7;
};
__device__ __forceinline__ Ks2W ks2_load_w(const float* __restrict__ t /* &tab[k * 32 + lane] */, const int rowStride, const float (&d)[8], const bool odd) {
    float v[8];
#pragma unroll
    for (int y = 0; y < 8; ++y) v[y] = d[y] * t[y * rowStride];
    Ks2W w;
    if (!odd) {
#pragma unroll
        for (int p = 0; p < 4; ++p) w.p[p] = ks_pk(v[2 * p], v[2 * p + 1]);
        w.s0 = 0.0f; w.s7 = 0.0f;
    } else {
#pragma unroll
        for (int p = 0; p < 3; ++p) w.p[p] = ks_pk(v[2 * p + 1], v[2 * p + 2]);
        w.p[3] = 0ull; w.s0 = v[0]; w.s7 = v[7];
    }
    return w;
}
__device__ __forceinline__ float ks_lds_again(const float* p) {        // a second, independent copy of a shared-memory word
    float v; asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(v) : "r"((unsigned)__cvta_generic_to_shared(p))); return v;
}
__device__ __forceinline__ Ks2G ks2_make_g(const float* __restrict__ t /* &tab[j * 32 + lane] */, const int rowStride) {
    // rows 1..6 are needed in two register pairings, (0,1)(2,3)(4,5)(6,7) and (1,2)(3,4)(5,6). A 64-bit operand is an aligned register
    // pair, so the second pairing needs its own copies; left to ptxas these become ~3 MOVs per use (it re-creates the pairs),
    // loaded a second time they cost 6 LDS per column offset and no MOV.
    float g[8], h[6];
#pragma unroll
    for (int y = 0; y < 8; ++y) g[y] = RTD_KS2_EARLY_G ? ks_lds_again(t + y * rowStride) : t[y * rowStride];
#pragma unroll
    for (int y = 0; y < 6; ++y) h[y] = ks_lds_again(t + (y + 1) * rowStride);
    Ks2G r;
#pragma unroll
    for (int p = 0; p < 4; ++p) r.e[p] = ks_pk(g[2 * p], g[2 * p + 1]);
#pragma unroll
    for (int p = 0; p < 3; ++p) r.o[p] = ks_pk(h[2 * p], h[2 * p + 1]);
    r.s0 = g[0]; r.s7 = g[7];
    return r;
}
// acc (window pairs) += g * w for the weight whose row-0 output has window index w0 (compile-time after unrolling)
template <int NP>
__device__ __forceinline__ void ks2_fma8(ks_u64 (&acc)[NP], const Ks2G& g, const Ks2W& w, const int w0) {
    if ((w0 & 1) == 0) {
#pragma unroll
        for (int p = 0; p < 4; ++p) acc[(w0 >> 1) + p] = ks_fma2(g.e[p], w.p[p], acc[(w0 >> 1) + p]);
    } else {
        ks_fma_hi(acc[w0 >> 1], g.s0, w.s0);
#pragma unroll
        for (int p = 0; p < 3; ++p) acc[((w0 + 1) >> 1) + p] = ks_fma2(g.o[p], w.p[p], acc[((w0 + 1) >> 1) + p]);
        ks_fma_lo(acc[(w0 + 7) >> 1], g.s7, w.s7);
    }
}
// Adds NV 128-bit words of register-held window pairs onto the patch column at `col` (16-byte aligned); window entries outside the
// output rows are zero and leave their patch rows unchanged.
template <int NV>
__device__ __forceinline__ void ks2_rmw(float* __restrict__ col, const ks_u64 (&acc)[2 * NV]) {
    float4* c = reinterpret_cast<float4*>(col);
#pragma unroll
    for (int q = 0; q < NV; ++q) {
        float4 v = c[q];
        const ks_u64 lo = ks_add2(ks_pk(v.x, v.y), acc[2 * q]), hi = ks_add2(ks_pk(v.z, v.w), acc[2 * q + 1]);
        ks_upk(lo, v.x, v.y); ks_upk(hi, v.z, v.w);
        c[q] = v;
    }
}

// All 8 source rows of the tile, vertical offsets |i| <= RB, column offsets 0..jmax, into the patch of a kernel instantiated for radii
// <= RBMAX.
//   tab    : tab[(y*TS + i)*32 + lane] = e_y[i] of the lane's 8 voxels (TS = table row stride in entries)
//   d      : dose of the lane's 8 voxels
//   patch  : column-major patch of this warp
template <int RB, int RBMAX, int TS>
__device__ __forceinline__ void ks2_tile(const float* __restrict__ tab, const float (&d)[8], float* __restrict__ patch, const int lane, const int jmax) {
    constexpr int P = 8, W = 2 * RB + 1, NRP = P + 2 * RB, PC = ks2_col_pitch(RBMAX);
    constexpr int Y0 = RBMAX - RB;            // patch row of the first output row (tile row -RB)
    constexpr int A = Y0 & 3, YS = Y0 - A;    // 128-bit window [YS, YS + 4 NV) around the NRP output rows: window index = output row + A
    constexpr int NV = (A + NRP + 3) / 4;
    static_assert(YS + 4 * NV <= ks2_rows(RBMAX), "window stays inside the column");
    Ks2W e[RB + 1];                           // weight index k = |i|: offsets -k and +k have row-0 window indices RB -+ k + A, same parity
#pragma unroll
    for (int k = 0; k <= RB; ++k) e[k] = ks2_load_w(tab + k * 32 + lane, TS * 32, d, ((RB + k + A) & 1) != 0);
    float* const col0 = patch + (lane + RBMAX) * PC + YS;     // the lane's own column (offset j = 0)
    // the 14 shared-memory words of the next column offset are requested as soon as the FMA block has consumed this one's, so
    // that they arrive under the patch read-modify-write instead of stalling the head of the next iteration; at the largest radii
    // the register file has no room for them next to the accumulators (ptxas starts shuffling pairs): there they are loaded in place
    constexpr bool EARLY_G = RTD_KS2_EARLY_G && RB <= 12;
    Ks2G g = ks2_make_g(tab + lane, TS * 32);
#pragma unroll 1
    for (int j = 0; j <= jmax; ++j) {       // jmax = applied radius (<= RB: a radius may run in the body of a larger one)
        if (!EARLY_G && j) g = ks2_make_g(tab + j * 32 + lane, TS * 32);
        ks_u64 acc[2 * NV];
#pragma unroll
        for (int m = 0; m < 2 * NV; ++m) acc[m] = 0ull;
#pragma unroll
        for (int i = 0; i < W; ++i) ks2_fma8<2 * NV>(acc, g, e[(i < RB) ? (RB - i) : (i - RB)], i + A);
        if (EARLY_G) g = ks2_make_g(tab + min(j + 1, RB) * 32 + lane, TS * 32);
        // the horizontal weight is symmetric: the same output column goes to patch columns x+j and x-j. Within one phase
        // every lane owns a different column.
        ks2_rmw<NV>(col0 + j * PC, acc);
        __syncwarp();
        if (j) { ks2_rmw<NV>(col0 - j * PC, acc); __syncwarp(); }
    }
}

// Batch radii 17..32 (patch of the RBMAX = 24 or RBMAX = 32 kernel). The row weights of all 8 voxels of a lane no longer fit the register
// file at once (8 x 33), so the vertical offsets are split into two bands that run one after the other over all column
// offsets j: the central band |i| <= 16 (8 x 17 weights, 40 output rows - the register shape of ks2_tile<16>) and the two
// outer bands 17 <= |i| <= RB, which share their 8 x (RB - 16) weights and produce RB - 9 output rows above and RB - 9
// below the tile. Same FMA count as a single pass, 86 instead of 72 rows of read-modify-write per column offset.
template <int RB, int RBMAX>
__device__ __forceinline__ void ks2_tile_big(const float* __restrict__ tab, const float (&d)[8], float* __restrict__ patch, const int lane) {
    constexpr int P = 8, RC = 16, PC = ks2_col_pitch(RBMAX), TS = RB + 1;
    static_assert(RB > RC && RB <= RBMAX, "radius class");
    ks2_tile<RC, RBMAX, TS>(tab, d, patch, lane, RB);        // central band: output rows -RC .. P-1+RC
    {   // outer bands: i = RC+1 .. RB below the source row (output rows RC+1 .. P-1+RB) and above it (-RB .. P-1-RC-1)
        constexpr int N2 = RB - RC, NB = P + N2 - 1;
        constexpr int YU = RBMAX - RB, AU = YU & 3, NVU = (AU + NB + 3) / 4;               // upper band: first output row -RB
        constexpr int YL = RBMAX + RC + 1, AL = YL & 3, NVL = (AL + NB + 3) / 4;           // lower band: first output row RC+1
        static_assert(((AL ^ (N2 - 1 + AU)) & 1) == 0, "both bands pair the rows of a weight the same way");
        float* const col0 = patch + (lane + RBMAX) * PC;
        Ks2W e[N2];                           // weight index k <-> offset RC + 1 + k: lower band window k + AL, upper band N2 - 1 - k + AU
#pragma unroll
        for (int k = 0; k < N2; ++k) e[k] = ks2_load_w(tab + (RC + 1 + k) * 32 + lane, TS * 32, d, ((k + AL) & 1) != 0);
#pragma unroll 1
        for (int j = 0; j <= RB; ++j) {
            const Ks2G g = ks2_make_g(tab + j * 32 + lane, TS * 32);
            ks_u64 lo[2 * NVL], up[2 * NVU];
#pragma unroll
            for (int m = 0; m < 2 * NVL; ++m) lo[m] = 0ull;
#pragma unroll
            for (int m = 0; m < 2 * NVU; ++m) up[m] = 0ull;
#pragma unroll
            for (int k = 0; k < N2; ++k) {
                ks2_fma8<2 * NVL>(lo, g, e[k], k + AL);                  // output row y + (RC + 1 + k)
                ks2_fma8<2 * NVU>(up, g, e[k], N2 - 1 - k + AU);         // output row y - (RC + 1 + k)
            }
            ks2_rmw<NVU>(col0 + j * PC + (YU - AU), up);
            ks2_rmw<NVL>(col0 + j * PC + (YL - AL), lo);
            __syncwarp();
            if (j) {
                ks2_rmw<NVU>(col0 - j * PC + (YU - AU), up);
                ks2_rmw<NVL>(col0 - j * PC + (YL - AL), lo);
                __syncwarp();
            }
        }
    }
}

// The layers of an item (slice, tile) are split by applied radius: those with radius <= 16 are accumulated by the RBMAX = 8
// kernel (if the largest of them is <= 8) or the RBMAX = 16 kernel (9..16), those with radius 17..32 by the RBMAX = 24
// kernel (if the largest of them is <= 24: 45 KB of patch + table per warp, 5 warps per SM) or the RBMAX = 32 kernel
// (25..32), whose 63 KB per warp leave only 3 warps per SM - small radii would run badly there.
template <int RBMAX, int WARPS>
__global__ void __launch_bounds__(WARPS * 32)
ks2_kernel(const Ks2Params p, int* __restrict__ workCounter, int* __restrict__ itemsDone) {
    constexpr int RBMIN = (RBMAX == 8) ? 0 : (RBMAX == 16) ? 9 : (RBMAX == 24) ? 17 : 25;      // the item's largest radius
    constexpr bool BIG = RBMAX > 16;                  // which layers of an item: applied radius 17..32 or 0..16
    constexpr int LAYER_RMIN = BIG ? 17 : 0;
    constexpr int PC = ks2_col_pitch(RBMAX), ROWS = ks2_rows(RBMAX);
    extern __shared__ __align__(16) float smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* patch = smem + (size_t)warp * ks2_warp_floats(RBMAX);
    float* tab = patch + ks2_patch_floats(RBMAX);
    // layer list entry q (layer | rb << 8) lives in the pad of patch column q / 8
    auto llist = [&](int q) -> uint16_t& { return reinterpret_cast<uint16_t*>(patch + (q >> 3) * PC + ROWS)[q & 7]; };
    // The RBMAX = 24 / 32 kernels run only 5 / 3 warps per SM: they walk the list of items that have a layer with radius > 16
    // (big_items_kernel) instead of scanning the whole beam; the list is empty for shallow targets.
    constexpr int NCH = 1;
    const int nz = p.bd->nz, first = p.bd->bfi;
    if (nz <= 0) return;
    const int nBig = *p.bigCount;
    const int nItems = BIG ? nBig : nz * p.tiles;
    // Beams that reach radii > 16 spread their layers over many different radii: the per-radius bodies (~108 KB of code in
    // the RBMAX = 16 kernel) then miss the instruction caches, and running odd radii in the next even body pays
    // (C4 superposition 15.7 -> 13.3 ms); beams that stay <= 16 are better off with the exact body (C3 1.14 vs 1.16 ms).
    const int bodyStep = p.bodyStep ? p.bodyStep : (nBig > 0 ? 2 : 1);
    const int R = p.Rx * p.Ry;
    const int bevPitch = p.Rx + 64;
    // the atomic that hands out the next item is issued a whole item ahead: its L2 round trip (~6 % of the small-radius kernel's
    // samples when it sat at the head of the loop) is hidden under the current item
    int itNext = 0;
    if (RTD_KS2_ITEM_PREFETCH && lane == 0) itNext = atomicAdd(workCounter, 1);
    for (;;) {
        if (!RTD_KS2_ITEM_PREFETCH && lane == 0) itNext = atomicAdd(workCounter, 1);
        int it = __shfl_sync(0xffffffffu, itNext, 0);
        if (it >= nItems * NCH) break;
        if (RTD_KS2_ITEM_PREFETCH && lane == 0) itNext = atomicAdd(workCounter, 1);
        const int idx = BIG ? (int)p.bigItems[it] : nItems - 1 - it / NCH;      // deepest slices (largest radii) first
        const int chunk = it % NCH;
        const int kk = idx / p.tiles, tile = idx - kk * p.tiles;
        // compact list of the layers that are live at this (slice, tile), with their applied radius
        int imax = -1, n = 0;
        for (int l0 = 0; l0 < p.L; l0 += 32) {
            const int l = l0 + lane;
            int rb = -1;
            if (l < p.L && kk < p.layerFirstPassive[l] - first) {
                const unsigned own = p.radMap[((size_t)l * nz + kk) * p.tiles + tile];
                if (own <= 32u) { rb = p.batchRad[l * 34 + own]; if (rb > KS2_MAXRB || (BIG ? rb <= 16 : rb > 16)) rb = -1; }   // > 32 is reported as an error by the host
            }
            const unsigned m = __ballot_sync(0xffffffffu, rb >= 0);
            const int pos = n + __popc(m & ((1u << lane) - 1u));       // ordinal among the accepted layers of the item
            if (rb >= 0 && pos % NCH == chunk) { llist(pos / NCH) = (uint16_t)(l | (rb << 8)); imax = max(imax, rb); }
            n += __popc(m);
        }
        n = (n > chunk) ? (n - chunk + NCH - 1) / NCH : 0;             // entries of this work unit
        imax = __reduce_max_sync(0xffffffffu, imax);
        if (imax < RBMIN || imax > RBMAX) continue;
        if (lane == 0) atomicAdd(itemsDone, 1);
        // zero the columns this item can touch (rows 0..ROWS-1 only: the pads hold the layer list)
        // (the column range is widened to whole groups of 4: the final reduction below adds 16-byte groups of 4 columns)
        const int nr = 8 + 2 * imax, c0 = RBMAX - imax;
        const int cz0 = c0 & ~3, ncz = ((c0 + 32 + 2 * imax + 3) & ~3) - cz0;
        for (int i = lane; i < ncz * (ROWS / 4); i += 32) {
            const int c = i / (ROWS / 4), q = i - c * (ROWS / 4);
            reinterpret_cast<float4*>(patch + (cz0 + c) * PC)[q] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        }
        __syncwarp();
        const int tx = tile % p.tilesX, ty = tile / p.tilesX;
        const size_t tileOff = (size_t)kk * R + (size_t)(ty * 8) * p.Rx + tx * 32 + lane;
        const size_t layerStride = (size_t)nz * R;
        // the next layer's 16 row segments are requested into L2 while the current layer is processed (a register prefetch would
        // cost 16 of the 255 registers the packed FMA block needs: ptxas then re-creates its operand pairs with MOVs)
        auto prefetchLayer = [&](const int q) {       // called by lanes 0..3: lane -> sector lane of each 128-byte row segment
            const size_t src = (size_t)(llist(q) & 0xff) * layerStride + tileOff + 7 * lane;
#pragma unroll
            for (int y = 0; y < 8; ++y) {
                asm volatile("prefetch.global.L2 [%0];" :: "l"(p.idd + src + (size_t)y * p.Rx));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(p.rsig + src + (size_t)y * p.Rx));
            }
        };
        for (int q = 0; q < n; ++q) {
            const int rb = llist(q) >> 8;
            // body = radius whose unrolled code runs: with bodyStep 2 / 4 a radius <= 16 uses the body of the next multiple
            // with zero weights beyond rb and the column loop still ending at rb - the same sums bit for bit, a few more
            // FMAs, but 1/2 or 1/4 of the code working set (instruction-fetch stalls, profiles/r01_ncu_full_ks2_c4.txt)
            const int body = BIG ? rb : min((rb + bodyStep - 1) & -bodyStep, RBMAX);
            float d[8], r[8];
            {
                const size_t src = (size_t)(llist(q) & 0xff) * layerStride + tileOff;
#pragma unroll
                for (int y = 0; y < 8; ++y) { d[y] = __ldg(p.idd + src + (size_t)y * p.Rx); r[y] = __ldg(p.rsig + src + (size_t)y * p.Rx); }
            }
            if (q + 1 < n && lane < 4) prefetchLayer(q + 1);      // a 32-ray row segment = 128 B = 4 sectors: one lane per sector
            bool any = false;
#pragma unroll
            for (int y = 0; y < 8; ++y) any |= d[y] > 0.0f;
            if (!__any_sync(0xffffffffu, any)) continue;      // __syncthreads_or(dose > 0), cuh:456
            // erf-difference weights (cuh:459-467) -> table: e[0] = erf(r/2), e[i] = (erf(r(i+1/2)) - erf(r(i-1/2)))/2
            // the 8 rows advance together so that 8 independent polynomial chains are in flight per lane; the argument
            // r (i + 1/2) is advanced by repeated addition
#if RTD_KS2_PACKED_ERF && RTD_KS2_R0_FAST
            // Applied radius 0 - half of all (tile, layer) pairs of a head plan, the entrance region of every layer: the voxel keeps its
            // dose, scaled by erf(r / 2)^2. Same operations as the general body at RB = 0 (table entry 2 h, weight d * t, FFMA2 t * w + 0,
            // 128-bit read-modify-write of the lane's own column) without the round trip through the weight table.
            if (!BIG && rb == 0 && body == 0) {
                ks_u64 acc[4];
#pragma unroll
                for (int pp = 0; pp < 4; ++pp) {
                    const ks_u64 h = ks_half_erf_pos2(ks_mul2(ks_pk(r[2 * pp], r[2 * pp + 1]), ks_bc(0.5f)));
                    float t0, t1; ks_upk(ks_add2(h, h), t0, t1);
                    acc[pp] = ks_fma2(ks_pk(t0, t1), ks_pk(d[2 * pp] * t0, d[2 * pp + 1] * t1), 0ull);
                }
                ks2_rmw<2>(patch + (lane + RBMAX) * PC + RBMAX, acc);
                __syncwarp();
                continue;
            }
#endif
#if RTD_KS2_PACKED_ERF
            {
                ks_u64 x[4], rr[4], hOld[4];
                float* t = tab + lane;
                const int rowStride = (body + 1) * 32;
#pragma unroll
                for (int pp = 0; pp < 4; ++pp) {
                    rr[pp] = ks_pk(r[2 * pp], r[2 * pp + 1]);
                    x[pp] = ks_mul2(rr[pp], ks_bc(0.5f));
                    hOld[pp] = ks_half_erf_pos2(x[pp]);
                    float a, b; ks_upk(ks_add2(hOld[pp], hOld[pp]), a, b);
                    t[(2 * pp) * rowStride] = a; t[(2 * pp + 1) * rowStride] = b;
                }
#pragma unroll 1
                for (int i = 1; i <= rb; ++i) {
                    t += 32;
#pragma unroll
                    for (int pp = 0; pp < 4; ++pp) {
                        x[pp] = ks_add2(x[pp], rr[pp]);
                        const ks_u64 hNew = ks_half_erf_pos2(x[pp]);
                        float a, b; ks_upk(ks_fma2(hOld[pp], ks_bc(-1.0f), hNew), a, b);
                        t[(2 * pp) * rowStride] = a; t[(2 * pp + 1) * rowStride] = b;
                        hOld[pp] = hNew;
                    }
                }
                for (int i = rb + 1; i <= body; ++i) {
                    t += 32;
#pragma unroll
                    for (int y = 0; y < 8; ++y) t[y * rowStride] = 0.0f;
                }
            }
#else
            {
                float x[8], erfOld[8];
                float* t = tab + lane;
                const int rowStride = (body + 1) * 32;
#pragma unroll
                for (int y = 0; y < 8; ++y) { x[y] = r[y] * 0.5f; erfOld[y] = ks_erf_pos(x[y]); t[y * rowStride] = erfOld[y]; }
#pragma unroll 1
                for (int i = 1; i <= rb; ++i) {
                    t += 32;
#pragma unroll
                    for (int y = 0; y < 8; ++y) {
                        x[y] += r[y];
                        const float erfNew = ks_erf_pos(x[y]);
                        t[y * rowStride] = 0.5f * (erfNew - erfOld[y]);
                        erfOld[y] = erfNew;
                    }
                }
                for (int i = rb + 1; i <= body; ++i) {
                    t += 32;
#pragma unroll
                    for (int y = 0; y < 8; ++y) t[y * rowStride] = 0.0f;
                }
            }
#endif
            __syncwarp();
            switch (body) {
#define RTD_KS2_CASE(N) case N: if constexpr (N <= RBMAX && N >= LAYER_RMIN) { if constexpr (N <= 16) ks2_tile<N, RBMAX, N + 1>(tab, d, patch, lane, rb); else ks2_tile_big<(N > 16) ? N : 17, BIG ? RBMAX : 32>(tab, d, patch, lane); } break;
                RTD_KS2_CASE(0) RTD_KS2_CASE(1) RTD_KS2_CASE(2) RTD_KS2_CASE(3) RTD_KS2_CASE(4) RTD_KS2_CASE(5) RTD_KS2_CASE(6)
                RTD_KS2_CASE(7) RTD_KS2_CASE(8) RTD_KS2_CASE(9) RTD_KS2_CASE(10) RTD_KS2_CASE(11) RTD_KS2_CASE(12) RTD_KS2_CASE(13)
                RTD_KS2_CASE(14) RTD_KS2_CASE(15) RTD_KS2_CASE(16) RTD_KS2_CASE(17) RTD_KS2_CASE(18) RTD_KS2_CASE(19) RTD_KS2_CASE(20)
                RTD_KS2_CASE(21) RTD_KS2_CASE(22) RTD_KS2_CASE(23) RTD_KS2_CASE(24) RTD_KS2_CASE(25) RTD_KS2_CASE(26) RTD_KS2_CASE(27)
                RTD_KS2_CASE(28) RTD_KS2_CASE(29) RTD_KS2_CASE(30) RTD_KS2_CASE(31) RTD_KS2_CASE(32)
#undef RTD_KS2_CASE
                default: break;
            }
            __syncwarp();
        }
        // one reduction of the touched part of the patch into the haloed BEV dose (z index = absolute step, 32-px halo):
        // patch (column pc, row pr)  ->  BEV (tile column - RBMAX + pc, tile row - RBMAX + pr). A lane takes one row of a 16-byte
        // group of 4 columns: lanes 0..15 / 16..31 = 16 consecutive rows of two adjacent groups, so that the four scalar reads are
        // conflict-free (consecutive rows of one patch column) and the two groups of a row fill one 32-byte sector of the BEV;
        // RBMAX, the BEV pitch and the tile origin are multiples of 4, so every group is 16-byte aligned: red.global.add.v4.f32
        float* out = p.bev + ((size_t)(kk + first) * (p.Ry + 64) + (ty * 8 + 32 - RBMAX)) * bevPitch + tx * 32 + 32 - RBMAX;
        for (int r0 = c0; r0 < c0 + nr; r0 += 16) {
            const int pr = r0 + (lane & 15);
            for (int g0 = cz0; g0 < cz0 + ncz; g0 += 8) {
                const int pc = g0 + 4 * (lane >> 4);
                if (pr < c0 + nr && pc < cz0 + ncz) {
                    const float* src = patch + pc * PC + pr;
                    const float4 v = make_float4(src[0], src[PC], src[2 * PC], src[3 * PC]);
                    if (v.x != 0.0f || v.y != 0.0f || v.z != 0.0f || v.w != 0.0f)
                        atomicAdd(reinterpret_cast<float4*>(out + (size_t)pr * bevPitch + pc), v);
                }
            }
        }
        __syncwarp();
    }
}

}  // namespace rtd
