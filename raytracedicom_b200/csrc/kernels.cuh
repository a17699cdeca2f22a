// sm_100a kernels of the proton pencil-beam dose path. New code; the arithmetic of every quantity that feeds
// an integer output of the reference (WEPL, sigma recurrence, ray weights, tile radius) is pinned with
// explicit PTX so it evaluates the reference's formulas with the same roundings as the reference's own
// compiled kernels (checked against its PTX: mul/add/fma .ftz, div.approx, sqrt.approx, lg2/ex2.approx).
//
// Reference formulas restated (all under /root/reference/src):
//   K1  trace_bev_kernel      <- fillBevDensityAndSp kernel_wrapper.cu:130-187, density_and_sp_tracer_params.cu:15-36
//       slice_reduce_kernel   <- sliceMinVar/sliceMaxVar kernel_wrapper.cuh:173-244 (launches cu:781-790)
//   K3a spot_to_ray_kernel    <- xConvGathResampGpu / yConvGathResampGpu gpu_convolution_2d.cu:16-59
//   K2  idd_sigma_kernel      <- fillIddAndSigma kernel_wrapper.cu:189-379, fill_idd_and_sigma_params.cu:28-83
//       tile_radius_kernel    <- tileRadCalc kernel_wrapper.cuh:256-313 ; batch_kernel <- host batching cu:966-976
//   K3b ks2_kernel (ks2.cuh)  <- kernelSuperposition<rad> kernel_wrapper.cuh:432-489
//   K4  bev_to_grid_kernel    <- primTransfDiv kernel_wrapper.cu:69-97, transfer_param_struct_div3.cu:22-34
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "host_geom.hpp"

namespace rtd {

// ---------------------------------------------------------------------------------------------------------
// pinned fp32 arithmetic
__device__ __forceinline__ float f_add(float a, float b) { float r; asm("add.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float f_sub(float a, float b) { float r; asm("sub.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float f_mul(float a, float b) { float r; asm("mul.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float f_fma(float a, float b, float c) { float r; asm("fma.rn.ftz.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c)); return r; }
__device__ __forceinline__ float f_div(float a, float b) { float r; asm("div.approx.ftz.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ float f_sqrt(float a) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ float f_rsqrt(float a) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ float f_lg2(float a) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ float f_ex2(float a) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ float f_neg(float a) { float r; asm("neg.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ float f_abs(float a) { float r; asm("abs.ftz.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ float f_ceil(float a) { float r; asm("cvt.rpi.ftz.f32.f32 %0, %1;" : "=f"(r) : "f"(a)); return r; }
__device__ __forceinline__ int f_toint_rz(float a) { int r; asm("cvt.rzi.ftz.s32.f32 %0, %1;" : "=r"(r) : "f"(a)); return r; }
__device__ __forceinline__ float u2f(unsigned a) { float r; asm("cvt.rn.f32.u32 %0, %1;" : "=f"(r) : "r"(a)); return r; }
__device__ __forceinline__ float i2f(int a) { float r; asm("cvt.rn.f32.s32 %0, %1;" : "=f"(r) : "r"(a)); return r; }

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ int warp_max_i(int v) { return __reduce_max_sync(0xffffffffu, v); }
__device__ __forceinline__ int warp_min_i(int v) { return __reduce_min_sync(0xffffffffu, v); }

// ---------------------------------------------------------------------------------------------------------
// K1: fan-beam density / WEPL tracer.
// One CTA = 32 x-adjacent rays (one warp width, so CT and LUT texture fetches of a warp are spatially coherent)
// x 9 warps: warps 0..7 each fetch TR_CH consecutive depth samples per round (8 independent tex3D + 16 tex1D
// in flight per thread), stage stopping power and HU in shared memory, and warp 8 performs the strictly
// sequential WEPL / entry / exit scan of the previous round (the reference's per-thread serial loop,
// cu:151-184, needs exactly this summation order for bit-identical WEPL). Position is advanced by repeated
// addition like the reference (cu:183) so sample positions are bit-identical too.
constexpr int TR_CH = 8;
constexpr int TR_NW = 8;
constexpr int TR_ROUND = TR_CH * TR_NW;
constexpr int TR_THREADS = (TR_NW + 1) * 32;

__global__ void __launch_bounds__(TR_THREADS)
trace_bev_kernel(const TraceParams p, cudaTextureObject_t ctTex, cudaTextureObject_t densTex, cudaTextureObject_t spTex,
                 cudaTextureObject_t densSpTex /* 0, or both LUTs as one float2 texture (same length and scale) */,
                 float* __restrict__ density, float* __restrict__ wepl, int* __restrict__ firstInside,
                 int* __restrict__ firstOutside, const int Rx, const int R,
                 const int* __restrict__ gate /* nullptr, or: run only if *gate != 0 (the fall-back behind trace_brick_kernel) */) {
    if (gate && *gate == 0) return;
    __shared__ float sSp[2][TR_ROUND][32];
    __shared__ float sHu[2][TR_ROUND][32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int ray = blockIdx.x * 32 + lane;
    const int ix = ray % Rx, iy = ray / Rx;
    const int S = (int)p.steps;
    const int nRounds = (S + TR_ROUND - 1) / TR_ROUND;
    const float fi = i2f(ix), fj = i2f(iy);

    if (warp < TR_NW) {
        // getStart (density_and_sp_tracer_params.cu:15) + HALF, getInc (cu:17)
        const float s1 = f_sub(1.0f, f_div(p.corner.z, p.distX));
        const float s2 = f_sub(1.0f, f_div(p.corner.z, p.distY));
        float px = f_add(f_add(p.transl.x, f_fma(f_mul(p.coefIdxI.x, fi), s1, f_mul(f_mul(p.coefIdxJ.x, fj), s2))), 0.5f);
        float py = f_add(f_add(p.transl.y, f_fma(f_mul(p.coefIdxI.y, fi), s1, f_mul(f_mul(p.coefIdxJ.y, fj), s2))), 0.5f);
        float pz = f_add(f_add(f_fma(f_mul(p.coefIdxI.z, fi), s1, f_mul(f_mul(p.coefIdxJ.z, fj), s2)), p.transl.z), 0.5f);
        const float dx = f_mul(f_sub(f_sub(p.coefOffset.x, f_div(f_mul(p.coefIdxI.x, fi), p.distX)), f_div(f_mul(p.coefIdxJ.x, fj), p.distY)), p.delta.z);
        const float dy = f_mul(f_sub(f_sub(p.coefOffset.y, f_div(f_mul(p.coefIdxI.y, fi), p.distX)), f_div(f_mul(p.coefIdxJ.y, fj), p.distY)), p.delta.z);
        const float dz = f_mul(f_sub(f_sub(p.coefOffset.z, f_div(f_mul(p.coefIdxI.z, fi), p.distX)), f_div(f_mul(p.coefIdxJ.z, fj), p.distY)), p.delta.z);
        for (int i = 0; i < warp * TR_CH; ++i) { px = f_add(dx, px); py = f_add(dy, py); pz = f_add(dz, pz); }
        for (int r = 0; r <= nRounds; ++r) {
            if (r < nRounds) {
                const int base = r * TR_ROUND + warp * TR_CH;
                float hu[TR_CH];
#pragma unroll
                for (int j = 0; j < TR_CH; ++j) {
                    hu[j] = tex3D<float>(ctTex, px, py, pz);
                    px = f_add(dx, px); py = f_add(dy, py); pz = f_add(dz, pz);
                }
#pragma unroll
                for (int j = 0; j < TR_CH; ++j) {
                    // the tracer is bound by the texture pipe (3 fetches per sample): when the density and stopping-power
                    // LUTs share length and scale they are one float2 texture, filtered per channel with the same weights
                    float d, s;
                    if (densSpTex) {
                        const float2 ds = tex1D<float2>(densSpTex, f_fma(hu[j], p.densityScale, 0.5f));
                        d = ds.x; s = ds.y;
                    } else {
                        d = tex1D<float>(densTex, f_fma(hu[j], p.densityScale, 0.5f));
                        s = tex1D<float>(spTex, f_fma(hu[j], p.spScale, 0.5f));
                    }
                    if (base + j < S) density[(size_t)(base + j) * R + ray] = d;
                    sSp[r & 1][warp * TR_CH + j][lane] = s;
                    sHu[r & 1][warp * TR_CH + j][lane] = hu[j];
                }
                for (int i = 0; i < TR_ROUND - TR_CH; ++i) { px = f_add(dx, px); py = f_add(dy, py); pz = f_add(dz, pz); }
            }
            __syncthreads();
        }
    } else {
        // stepLen (density_and_sp_tracer_params.cu:32-36)
        const float ddx = f_div(f_fma(p.delta.x, fi, p.corner.x), p.distX);
        const float ddy = f_div(f_fma(p.delta.y, fj, p.corner.y), p.distY);
        const float stepLen = f_mul(f_abs(p.delta.z), f_sqrt(f_fma(ddy, ddy, f_fma(ddx, ddx, 1.0f))));
        float cumSp = 0.0f, cumHu = 0.0f;
        int before = -1, last = -1;
        for (int r = 0; r <= nRounds; ++r) {
            if (r > 0) {
                const int base = (r - 1) * TR_ROUND;
                const int n = min(TR_ROUND, S - base);
#pragma unroll 8
                for (int s = 0; s < n; ++s) {
                    const float hu = sHu[(r - 1) & 1][s][lane];
                    const float sp = sSp[(r - 1) & 1][s][lane];
                    cumHu = f_add(cumHu, hu);
                    cumSp = f_fma(stepLen, sp, cumSp);
                    if (cumHu < 150.0f) before = base + s;
                    if (hu > 150.0f) last = base + s;
                    wepl[(size_t)(base + s) * R + ray] = cumSp;
                }
            }
            __syncthreads();
        }
        firstInside[ray] = before + 1;
        firstOutside[ray] = last + 1;
    }
}

// Per-depth-slice minimum WEPL plus the beam-wide entry / exit steps (three reference launches + three D2H
// copies, cu:781-790, become one launch and one copy). grid = steps, block = 256. out: [0]=min firstInside,
// [1]=max firstOutside, weplMin[steps].
__global__ void __launch_bounds__(256)
slice_reduce_kernel(const float* __restrict__ wepl, const int* __restrict__ firstInside, const int* __restrict__ firstOutside,
                    const int R, float* __restrict__ weplMin, int* __restrict__ beamFirstInside, int* __restrict__ beamFirstOutside) {
    __shared__ float sMin[8];
    __shared__ int sA[8], sB[8];
    const int k = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float4* row = reinterpret_cast<const float4*>(wepl + (size_t)k * R);
    float m = __int_as_float(0x7f800000);
    for (int i = threadIdx.x; i < R / 4; i += 256) { const float4 v = row[i]; m = fminf(fminf(m, v.x), fminf(v.y, fminf(v.z, v.w))); }
    m = warp_min(m);
    if (lane == 0) sMin[warp] = m;
    if (k == 0) {
        int a = 0x7fffffff, b = -0x7fffffff;
        for (int i = threadIdx.x; i < R; i += 256) { a = min(a, firstInside[i]); b = max(b, firstOutside[i]); }
        a = warp_min_i(a); b = warp_max_i(b);
        if (lane == 0) { sA[warp] = a; sB[warp] = b; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        float mm = sMin[0];
        for (int i = 1; i < 8; ++i) mm = fminf(mm, sMin[i]);
        weplMin[k] = mm;
        if (k == 0) {
            int a = sA[0], b = sB[0];
            for (int i = 1; i < 8; ++i) { a = min(a, sA[i]); b = max(b, sB[i]); }
            *beamFirstInside = a; *beamFirstOutside = b;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Per-beam set-up ON THE DEVICE. The reference (and round 1 of this engine) brings the tracer's reductions to the host, derives
// the cut-off steps, the per-layer parameters and the transfer box there (kernel_wrapper.cu:792-849, 922-925, 1185-1218) and only
// then launches the rest of the beam: a host round trip in the middle of every beam. Here one small CTA does that arithmetic and
// leaves the results in device memory (BeamDev, LayerParams[], entry sigmas, step table); every later kernel of the beam reads its
// data-dependent parameters from there and is launched with a grid that depends on the geometry alone, so a plan is issued without
// any host synchronisation and the host reads the integers back after the final sync for its reports.
// Host formulas are evaluated with IEEE round-to-nearest intrinsics in the host's operation order (the host code is compiled
// without FMA contraction): the same bits as round 1's host code, which the integer parity tests pin.
struct LayerParams {          // per-layer constants, kernel_wrapper.cu:829-849, 922-925
    float energyIdx, energyScale, peakDepth;
    int afterLast;
    int order;                // launch order: grid.z = i runs layer layers[i].order, deepest layers (longest depth scans) first
};
struct LayerIn {              // what the host knows about a layer before the tracer has run
    float energyIdx, energyScale, peakDepth, thr /* BP_DEPTH_CUTOFF x peakDepth */, ca, cb /* sigmaSqAirCoefs, host IEEE */, sigX, sigY;
    float sigScale;           // 0.97 for NUCLEAR_CORR == GAUSS_FIT (kernel_wrapper.cu:842-847), else 1
};
struct GridParams {
    TransferParams t;
    int minX, minY, minZ, maxX, maxY, maxZ;   // inclusive bbox in dose indices
    int nx, ny, nz;                           // dose dims
    int W, H, first, zEnd;                    // BEV dims (haloed) and valid slice range [first, zEnd)
    int atomic;                               // 1: red.global.add.f32 (the dose grid is shared by several GPUs over NVLink)
    int tilesX, tilesY, zChunks;              // work units of bev_to_grid_kernel: 32 x 8 columns x a chunk of the z range; 0 = nothing to do
};
static_assert(sizeof(GridParams) % 4 == 0 && sizeof(GridParams) / 4 <= 256, "GridParams is copied by one CTA");
struct BeamDev {
    int bfi, bfo, Zg, nz;                     // beamFirstInside / Outside, beamFirstGuaranteedPassive, slices kept per layer = max(Zg - bfi, 0)
    float entryZ, pixelSpX, pixelSpY, pad0;   // K3a: ray pitch at the entry plane
    GridParams grid;                          // K4
};
struct SetupIn {
    int S, L;
    float rayResZ, rayOffZ, distX, distY, rayDx, rayDy;
    float thrAll;                             // BP_DEPTH_CUTOFF x peak depth of the beam's highest energy
    IdxT fitf; Affine gtdi; float dist[2];    // rayIdxToGantry, gantryToDoseIdx, source distances: transfer box
    TransferParams t0; float ftfiOffZ;        // transfer parameters for shift (32, 32, 0); the z shift -bfi is added here
    int Rx, Ry, nx, ny, nz, smCount, atomic;
    FillGeom g;                               // step table
};

namespace ieee {   // host arithmetic on the device: one rounding per operation, no contraction, IEEE division and square root
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return add(add(mul(a.x, b.x), mul(a.y, b.y)), mul(a.z, b.z)); }
// float3_from_fan_transform.cu:37-42 as host_geom.hpp's fromFan evaluates it
__device__ __forceinline__ V3 fromFan(const IdxT& fitf, const float distX, const float distY, const Affine& gtii, V3 f) {
    V3 p; p.x = add(mul(f.x, fitf.delta.x), fitf.offset.x); p.y = add(mul(f.y, fitf.delta.y), fitf.offset.y); p.z = add(mul(f.z, fitf.delta.z), fitf.offset.z);
    p.x = mul(p.x, sub(1.0f, div(p.z, distX)));
    p.y = mul(p.y, sub(1.0f, div(p.z, distY)));
    V3 r; r.x = add(dot3(gtii.m.r0, p), gtii.v.x); r.y = add(dot3(gtii.m.r1, p), gtii.v.y); r.z = add(dot3(gtii.m.r2, p), gtii.v.z);
    return r;
}
__device__ __forceinline__ int firstLargerOrdered(const float* __restrict__ list, int n, float value) {     // vector_find.h:62-82
    int upper = n - 1, lower = 0;
    if (list[n - 1] <= value) return upper;
    if (list[0] > value) return 0;
    while (upper - lower > 1) { const int pivot = (upper + lower) / 2; if (list[pivot] <= value) lower = pivot; else upper = pivot; }
    return lower + 1;
}
}  // namespace ieee

// block = 256 threads, one CTA per beam. trace: [0] = beamFirstInside, [1] = beamFirstOutside (ints), [2..] weplMin[S] (slice_reduce_kernel).
__global__ void __launch_bounds__(256)
beam_setup_kernel(const SetupIn in, const int* __restrict__ trace, const LayerIn* __restrict__ lin,
                  BeamDev* __restrict__ bd, LayerParams* __restrict__ layers, float2* __restrict__ sigmas, float2* __restrict__ stepTab) {
    __shared__ int sZg, sBfi;
    __shared__ int sAfter[256];
    const float* weplMin = reinterpret_cast<const float*>(trace + 2);
    if (threadIdx.x == 0) {
        sBfi = trace[0];
        sZg = min(ieee::firstLargerOrdered(weplMin, in.S, in.thrAll), trace[1]);      // kernel_wrapper.cu:792-796
    }
    __syncthreads();
    const int bfi = sBfi, Zg = sZg, nz = max(Zg - bfi, 0);
    const float entryZ = ieee::add(ieee::mul((float)bfi, in.rayResZ), in.rayOffZ);
    for (int l = threadIdx.x; l < in.L; l += 256) {
        const LayerIn li = lin[l];
        // entry sigma (cu:838-841): sqrt(ca z^2 + cb z + sigma^2), host arithmetic
        const float air = ieee::add(ieee::mul(ieee::mul(li.ca, entryZ), entryZ), ieee::mul(li.cb, entryZ));
        float2 es = make_float2(__fsqrt_rn(ieee::add(air, ieee::mul(li.sigX, li.sigX))), __fsqrt_rn(ieee::add(air, ieee::mul(li.sigY, li.sigY))));
        if (li.sigScale != 1.0f) { es.x = ieee::mul(li.sigScale, es.x); es.y = ieee::mul(li.sigScale, es.y); }
        sigmas[l] = es;
        const unsigned localAfterLast = (unsigned)ieee::firstLargerOrdered(weplMin, in.S, li.thr);     // cu:922-924
        sAfter[l] = (int)min(localAfterLast, (unsigned)Zg);
    }
    __syncthreads();
    for (int l = threadIdx.x; l < in.L; l += 256) {
        // launch order of the layers in the IDD / sigma kernel: longest depth scan first (stable)
        int rank = 0;
        for (int m = 0; m < in.L; ++m) rank += (sAfter[m] > sAfter[l]) || (sAfter[m] == sAfter[l] && m < l);
        const LayerIn li = lin[l];
        layers[l].energyIdx = li.energyIdx; layers[l].energyScale = li.energyScale; layers[l].peakDepth = li.peakDepth;
        layers[l].afterLast = sAfter[l];
        layers[rank].order = l;
    }
    // voxelWidth(k) (fill_idd_and_sigma_params.cu:42-46) and stepVol(k) (cu:72) depend on the step only: one table per beam,
    // evaluated with the instruction sequence the depth scan used to run per (ray, step) - bit-identical
    for (int kk = threadIdx.x; kk < nz; kk += 256) {
        const unsigned k = (unsigned)(bfi + kk);
        const float fk = u2f(k);
        const float zk = f_fma(in.g.delta.z, fk, in.g.corner.z);
        const float vwx = f_mul(in.g.delta.x, f_sub(1.0f, f_div(zk, in.g.distX)));
        const float vwy = f_mul(in.g.delta.y, f_sub(1.0f, f_div(zk, in.g.distY)));
        stepTab[kk] = make_float2(f_mul(f_add(vwx, vwy), 0.5f), f_fma(in.g.volSq, u2f(k * k), f_fma(in.g.volLin, fk, in.g.volConst)));
    }
    if (threadIdx.x == 0) {
        BeamDev b;
        b.bfi = bfi; b.bfo = trace[1]; b.Zg = Zg; b.nz = nz; b.entryZ = entryZ; b.pad0 = 0.0f;
        b.pixelSpX = ieee::mul(in.rayDx, ieee::sub(1.0f, ieee::div(entryZ, in.distX)));       // cu:849-854
        b.pixelSpY = ieee::mul(in.rayDy, ieee::sub(1.0f, ieee::div(entryZ, in.distY)));
        GridParams& gp = b.grid;
        gp.t = in.t0;
        gp.t.globalOffset.z = ieee::add(in.ftfiOffZ, -(float)bfi);
        gp.nx = in.nx; gp.ny = in.ny; gp.nz = in.nz;
        gp.W = in.Rx + 64; gp.H = in.Ry + 64; gp.first = bfi; gp.zEnd = Zg; gp.atomic = in.atomic;
        gp.tilesX = gp.tilesY = gp.zChunks = 0;
        gp.minX = gp.minY = gp.minZ = 0; gp.maxX = gp.maxY = gp.maxZ = -1;
        if (nz > 0) {
            // K4 box (cu:1185-1218); the z-range uses Zg instead of beamFirstCalculatedPassive: a superset box that only adds
            // voxels whose interpolated BEV dose is exactly 0
            const float xv[2] = {-32.0f, (float)(in.Rx + 32 - 1)}, yv[2] = {-32.0f, (float)(in.Ry + 32 - 1)}, zv[2] = {(float)bfi, (float)(Zg - 1)};
            V3 mx = v3(-1.f, -1.f, -1.f), mn = v3(100000.f, 100000.f, 100000.f);
            for (int c = 0; c < 8; ++c) {
                const V3 p = ieee::fromFan(in.fitf, in.dist[0], in.dist[1], in.gtdi, v3(xv[c & 1], yv[(c >> 1) & 1], zv[c >> 2]));
                mx.x = fmaxf(mx.x, p.x); mx.y = fmaxf(mx.y, p.y); mx.z = fmaxf(mx.z, p.z);
                mn.x = fminf(mn.x, p.x); mn.y = fminf(mn.y, p.y); mn.z = fminf(mn.z, p.z);
            }
            gp.minX = max(((int)floorf(mn.x) / 32) * 32, 0); gp.minY = max((int)floorf(mn.y), 0); gp.minZ = max((int)floorf(mn.z), 0);
            gp.maxX = min((int)ceilf(mx.x), in.nx - 1); gp.maxY = min((int)ceilf(mx.y), in.ny - 1); gp.maxZ = min((int)ceilf(mx.z), in.nz - 1);
            if (gp.maxX >= gp.minX && gp.maxY >= gp.minY && gp.maxZ >= gp.minZ) {
                gp.tilesX = (gp.maxX - gp.minX + 32) / 32; gp.tilesY = (gp.maxY - gp.minY + 8) / 8;
                const int tilesXY = gp.tilesX * gp.tilesY, nzBox = gp.maxZ - gp.minZ + 1;
                gp.zChunks = max(1, min(min((8 * in.smCount + tilesXY - 1) / tilesXY, nzBox), 64));     // ~8 work units per SM
            }
        }
        *bd = b;
    }
}

// ---------------------------------------------------------------------------------------------------------
// K3a: spot weights -> ray weights. One CTA = (32 ray columns, one layer): the x-pass result for all Sy spot
// rows stays in shared memory and the y-pass reads it from there (the reference round-trips an intermediate
// array through global memory between two launches, gpu_convolution_2d.cu:69-70).
struct ConvParams {           // the ray pitch at the entry plane (pixelSpX / Y) depends on the tracer: BeamDev
    int inW, inH, outW, outH;
    float offX, offY, deltaX, deltaY, cutoff;
    float rayWeightCutoff;    // RAY_WEIGHT_CUTOFF: a 32-ray row segment with no weight above it is dead for the depth scan
    int listAll;              // 1: every segment is listed (rtd_opts.fullIntermediates)
};

__device__ __forceinline__ float conv_weight(float dist, float rSig) {   // gpu_convolution_2d.cu:28
    return f_mul(f_sub(erff(f_mul(rSig, f_add(dist, 0.5f))), erff(f_mul(rSig, f_add(dist, -0.5f)))), 0.5f);
}

__global__ void __launch_bounds__(256)
spot_to_ray_kernel(const float* __restrict__ in, const float2* __restrict__ sigmas, float* __restrict__ out, const ConvParams c,
                   const BeamDev* __restrict__ bd, int* __restrict__ segCount /*[L]*/, uint16_t* __restrict__ segList /*[L][segs]*/) {
    extern __shared__ float sInterm[];   // [inH][32]
    if (bd->nz <= 0) return;             // the beam never enters the patient
    const float pixelSpX = bd->pixelSpX, pixelSpY = bd->pixelSpY;
    const int layer = blockIdx.y;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int ox = blockIdx.x * 32 + tx;
    const float2 sg = sigmas[layer];
    {   // x pass (gpu_convolution_2d.cu:16-35)
        const float sigEff = f_div(sg.x, pixelSpX);
        const float rSig = f_div(f_rsqrt(2.0f), sigEff);
        const float lim = f_fma(sigEff, c.cutoff, 0.5f);
        const float fo = i2f(ox);
        const int c0 = f_toint_rz(f_ceil(f_div(f_sub(f_sub(fo, lim), c.offX), c.deltaX)));
        for (int y = ty; y < c.inH; y += 8) {
            const float* rowp = in + ((size_t)layer * c.inH + y) * c.inW;
            float res = 0.0f;
            int ci = c0;
            float dist = f_sub(f_fma(c.deltaX, i2f(ci), c.offX), fo);
            while (dist < lim) {
                if (ci >= 0 && ci < c.inW) res = f_fma(rowp[ci], conv_weight(dist, rSig), res);
                ++ci;
                dist = f_sub(f_fma(c.deltaX, i2f(ci), c.offX), fo);
            }
            sInterm[y * 32 + tx] = res;
        }
    }
    __syncthreads();
    {   // y pass (gpu_convolution_2d.cu:37-59)
        const float sigEff = f_div(sg.y, pixelSpY);
        const float rSig = f_div(f_rsqrt(2.0f), sigEff);
        const float lim = f_fma(sigEff, c.cutoff, 0.5f);
        for (int oy = ty; oy < c.outH; oy += 8) {
            const float fo = i2f(oy);
            int ci = f_toint_rz(f_ceil(f_div(f_sub(f_sub(fo, lim), c.offY), c.deltaY)));
            float res = 0.0f;
            float dist = f_sub(f_fma(c.deltaY, i2f(ci), c.offY), fo);
            while (dist < lim) {
                if (ci >= 0 && ci < c.inH) res = f_fma(sInterm[ci * 32 + tx], conv_weight(dist, rSig), res);
                ++ci;
                dist = f_sub(f_fma(c.deltaY, i2f(ci), c.offY), fo);
            }
            out[((size_t)layer * c.outH + oy) * c.outW + ox] = res;
            // a warp holds one 32-ray row segment of the layer: the segments with a weight above the cut-off are the work items of
            // the depth scan (idd_sigma_kernel), listed here while their weights are in registers
            const bool liveSeg = __any_sync(0xffffffffu, !(res < c.rayWeightCutoff)) || c.listAll;
            if (tx == 0 && liveSeg) segList[(size_t)layer * (c.outH * gridDim.x) + atomicAdd(segCount + layer, 1)] = (uint16_t)(oy * gridDim.x + blockIdx.x);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// K2: integrated depth dose + lateral sigma for ALL energy layers in one launch (grid.z = layer; the
// reference launches once per layer and syncs with the host three times per layer, cu:916-976).
struct IddParams {
    FillGeom g;
    float rrlScale, bpCutoff, rayWeightCutoff;
    int Rx, Ry;
    int nozzle;               // 0 = the reference's -DNO_NOZZLE build (fill_idd_and_sigma_params.cu:76-82)
    int doseToMedium;         // 1 = the reference built without -DDOSE_TO_WATER (kernel_wrapper.cu:314-318)
    int nuclear;              // rtd_opts.nuclearCorr: 0 off, 1 SOUKUP, 2 FLUKA, 3 GAUSS_FIT (kernel_wrapper.cu:228-245, 300-302, 320-341)
    float eRefSq, sigmaDelta; // E_s^2 and the empirical widening of the model (198.81 / 0.21 without NUCLEAR_CORR)
    int writeAll;             // 1: every (ray, step) of idd / rSigmaEff is written (rtd_opts.fullIntermediates, for rtd_debug_fetch);
                              // 0: row segments without dose are left alone - the superposition never reads them (see the step lambda)
};

// Thread = (ray, layer), warp = 32-ray row segment. The depth scan of a segment is a serial chain whose length differs wildly
// between segments: rays outside the field never start (two thirds of all segments on C3), shallow layers stop early. A grid of
// one CTA per (segment, layer) leaves that to the block scheduler, which deals CTAs to SMs in launch order: ncu showed the busiest
// scheduler issuing 1.75 x the average (profiles/r02_ncu_full_c3_mid.txt) and the kernel lasting as long as that one. Here the
// pre-convolution lists the live segments of every layer (spot_to_ray_kernel), and persistent warps pull (layer, segment) items
// from ONE queue ordered by decreasing scan length (longest-processing-time first): every scheduler stays busy until the queue is
// empty. Dead segments are never visited: their row minima / firstPassive are the fill values beam_clear_kernel wrote.
// The per-step table the CTA used to build in shared memory is the beam's stepTab (beam_setup_kernel), read through L1.
#ifndef RTD_IDD_BLOCK
#define RTD_IDD_BLOCK 4
#endif
#ifndef RTD_IDD_PRELOAD
#define RTD_IDD_PRELOAD 0
#endif
constexpr int IDD_WARPS = 2;
__global__ void __launch_bounds__(IDD_WARPS * 32)
idd_sigma_kernel(const float* __restrict__ density, const float* __restrict__ wepl, const float* __restrict__ rayWeights,
                 const int* __restrict__ firstInside, const int* __restrict__ firstOutside,
                 const LayerParams* __restrict__ layers, const BeamDev* __restrict__ bd, const float2* __restrict__ stepTab, const IddParams q,
                 cudaTextureObject_t iddTex, cudaTextureObject_t rrlTex, cudaTextureObject_t nucTex,
                 float* __restrict__ idd, float* __restrict__ rsig, float* __restrict__ rowMin,
                 int* __restrict__ firstPassive, int* __restrict__ layerFirstPassive,
                 const int L, const int* __restrict__ segCount, const uint16_t* __restrict__ segList, int* __restrict__ queue) {
    const int nz = bd->nz;
    if (nz <= 0) return;
    __shared__ int sPrefix[257];          // items before the i-th layer in launch order (deepest first)
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int i = 0; i < L; ++i) { sPrefix[i] = acc; acc += segCount[layers[i].order]; }
        sPrefix[L] = acc;
    }
    __syncthreads();
    const int total = sPrefix[L];
    const int tilesX = q.Rx / 32;
    const int segsPerLayer = tilesX * q.Ry;
    const int R = q.Rx * q.Ry;
    const int lane = threadIdx.x & 31;
    const unsigned first = (unsigned)bd->bfi;
    const float inf = __int_as_float(0x7f800000);
  for (;;) {
    int item = 0;
    if (lane == 0) item = atomicAdd(queue, 1);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= total) break;
    int li = 0;
    while (sPrefix[li + 1] <= item) ++li;          // L <= 255 and the items of a warp mostly advance through the layers in order
    const int l = layers[li].order;
    const int seg = segList[(size_t)l * segsPerLayer + (item - sPrefix[li])];
    const int segX = seg % tilesX;
    const int x = segX * 32 + lane;
    const int y = seg / tilesX;
    const int ray = y * q.Rx + x;
    const LayerParams lp = layers[l];

    const int firstIn = firstInside[ray];
    unsigned afterLast = (unsigned)min(firstOutside[ray], lp.afterLast);
    const float rayWeight = rayWeights[(size_t)l * R + ray];
    bool live = true;
    if (rayWeight < q.rayWeightCutoff || afterLast < first) { live = false; afterLast = 0; }

    // initStepAndAirDiv (fill_idd_and_sigma_params.cu:28-40) with the device-side sigmaSqAirCoefs (cu:74-83)
    float ca = 0.0f, cb = 0.0f;
    if (q.nozzle) { ca = f_div(0.00270f, f_add(lp.peakDepth, -4.50f)); cb = f_div(-4.39f, f_add(lp.peakDepth, -3.86f)); }
    const float dzs = q.g.delta.z;
    const float airQuad = f_mul(dzs, f_mul(ca, dzs));
    const float airLin = f_fma(cb, dzs, f_mul(q.g.corner.z, f_mul(f_add(ca, ca), dzs)));
    const float stepLength = f_abs(dzs);

    float incScat = 0.0f, incinc = 0.0f;
    float incDiv = f_fma(airQuad, f_fma(u2f(first), 2.0f, -1.0f), airLin);
    float sigmaSq = f_neg(incDiv);
    float res = 0.0f, rs = 0.0f, wOld = 0.0f, dOld = 0.0f;
    const float thr = f_mul(lp.peakDepth, q.bpCutoff);

    // per-thread base pointers + one 32-bit running offset per array (a slab of one layer / one beam stays far below 2^32
    // elements): one IMAD.WIDE per access instead of 64-bit index chains
    const float* const weplP = wepl + (size_t)first * R + ray;
    const float* const densP = density + (size_t)first * R + ray;
    float* const iddP = idd + (size_t)l * nz * R + ray;
    float* const rsigP = rsig + (size_t)l * nz * R + ray;
    float* const rowMinP = rowMin + ((size_t)l * nz * q.Ry + y) * (q.Rx / 32) + segX;
    const unsigned rmStride = (unsigned)(q.Ry * (q.Rx / 32));
    unsigned off = 0, rmOff = 0;
    // The depth scan is a dependent chain per ray, and the load -> texture -> arithmetic latency of one step is what a step costs
    // unless the requests run ahead of their use: see the slots below. Coordinates and order of the arithmetic are the reference's.
    const unsigned last = (unsigned)lp.afterLast;
    const float eRow = f_add(lp.energyIdx, 0.5f);
    // The scattering power per unit 1/X0 of a step (t1: residual energy from the mean WEPL of the step, 4 of the step's 7 MUFU
    // operations) depends on WEPL alone, not on the running sums: it is evaluated one step ahead as well, which leaves
    // only the short sigma^2 recurrence on the loop-carried chain.
    auto scatPower = [&](const float wPrev, const float wCur) {
        const float resE = f_ex2(f_mul(f_lg2(f_sub(lp.peakDepth, f_mul(f_add(wPrev, wCur), 0.5f))), 0.5649718f));
        const float e1 = f_fma(resE, 8.639415f, 938.3f);
        const float betaP = f_sub(e1, f_div(880406.875f, e1));
        return f_mul(f_div(q.eRefSq, f_mul(betaP, betaP)), stepLength);
    };
#if RTD_IDD_BLOCK
    // Round 2, late: the scan in blocks of RTD_IDD_BLOCK steps. Everything of a step that does not depend on the running sums - the
    // two loads, the two (three with the nuclear option) LUT fetches, the scattering power with its four MUFU operations, i.e. nine
    // tenths of the step's instructions and all of its long latencies - is evaluated for the WHOLE block first, as independent
    // instruction streams the hardware overlaps; then the steps run in order with only the sigma^2 recurrence (four dependent adds),
    // the range test and the outputs. Same operations on the same operands in the same order per ray as the four-slot pipeline it
    // replaces (kept below, RTD_IDD_BLOCK 0), whose three steps of look-ahead still left 4.2 warp-cycles of long-scoreboard stall per
    // issued instruction (profiles/r02_ncu_full_c3.txt).
    constexpr int NB = RTD_IDD_BLOCK;
    unsigned k = first;
    float wPrev = 0.0f;                      // WEPL of the step before the block (0 in front of the first step, cu:266)
    // the WEPL / density of a block are requested one block ahead (RTD_IDD_PRELOAD): a block then starts with its LUT fetches
    float pw[NB], prho[NB];
#pragma unroll
    for (int s = 0; s < NB; ++s) {
        const bool on = RTD_IDD_PRELOAD && live && first + (unsigned)s < last;
        pw[s] = on ? weplP[(unsigned)s * (unsigned)R] : 0.0f;
        prho[s] = on ? densP[(unsigned)s * (unsigned)R] : 0.0f;
    }
    while (k < last && __any_sync(0xffffffffu, live)) {
        const int nb = min(NB, (int)(last - k));                 // warp-uniform
        float bw[NB], brho[NB], bd_[NB], bX[NB], bnw[NB]; float2 bst[NB];
        float wLast = wPrev;
#pragma unroll
        for (int s = 0; s < NB; ++s) {
            const bool on = live && s < nb;
            if (RTD_IDD_PRELOAD) { bw[s] = pw[s]; brho[s] = prho[s]; }
            else {
                bw[s] = on ? weplP[off + (unsigned)s * (unsigned)R] : 0.0f;
                brho[s] = on ? densP[off + (unsigned)s * (unsigned)R] : 0.0f;
            }
            bst[s] = (s < nb) ? __ldg(stepTab + (k + s - first)) : make_float2(0.0f, 0.0f);
        }
        if (RTD_IDD_PRELOAD) {
#pragma unroll
            for (int s = 0; s < NB; ++s) {
                const bool on = live && k + (unsigned)(NB + s) < last;
                pw[s] = on ? weplP[off + (unsigned)(NB + s) * (unsigned)R] : 0.0f;
                prho[s] = on ? densP[off + (unsigned)(NB + s) * (unsigned)R] : 0.0f;
            }
        }
#pragma unroll
        for (int s = 0; s < NB; ++s) {
            const bool on = live && s < nb;
            const float wBefore = s ? bw[s - 1] : wPrev;
            bd_[s] = 0.0f; bX[s] = 0.0f; bnw[s] = 0.0f;
            if (on) {
                bd_[s] = tex2D<float>(iddTex, f_fma(bw[s], lp.energyScale, 0.5f), eRow);
                const float x = tex1D<float>(rrlTex, f_fma(brho[s], q.rrlScale, 0.5f));
                const float t1 = scatPower(wBefore, bw[s]);
                const float rRl = f_mul(brho[s], x);
                const float thetaSq = f_mul(rRl, t1);
                bX[s] = f_mul(stepLength, f_fma(rRl, t1, thetaSq));
                if (q.nuclear) bnw[s] = tex2D<float>(nucTex, f_fma(f_mul(f_add(bw[s], wBefore), 0.5f), lp.energyScale, 0.5f), eRow);
            }
        }
#pragma unroll
        for (int s = 0; s < NB; ++s) {
            if (s < nb) {
                const unsigned kk = k + (unsigned)s;
                const float w = bw[s], rho = brho[s];
                wLast = w;
                if (live) {
                    if (w < lp.peakDepth) {
                        sigmaSq = f_add(sigmaSq, f_add(incScat, incDiv));
                        incinc = f_fma(stepLength, bX[s], incinc);
                        incScat = f_add(incScat, incinc);
                        incDiv = f_fma(airQuad, 2.0f, incDiv);
                    } else if (q.nuclear != 3) {       // the empirical fix of the sigma dip after the peak is compiled out for GAUSS_FIT (cu:300-302)
                        sigmaSq = f_fma(f_mul(f_add(incScat, incDiv), -1.5f), rho, sigmaSq);
                    }
                    rs = f_div(bst[s].x, f_mul(f_add(f_sqrt(sigmaSq), q.sigmaDelta), 1.41421356f));
                    const bool within = !(w > thr);
                    live = within && (kk != afterLast);
                    afterLast = within ? afterLast : kk;
                    const float vol = bst[s].y;
                    const float mass = q.doseToMedium ? f_mul(rho, vol) : f_mul(f_sub(w, wOld), vol);
                    const float d = bd_[s];
                    if (mass > 1e-2f) {
                        // (nuclear option: the halo's share is taken out of the primary dose, cu:320-330; see the slot version below)
                        if (q.nuclear) res = f_div(f_mul(f_mul(f_sub(1.0f, bnw[s]), rayWeight), f_sub(d, dOld)), mass);
                        else res = f_div(f_mul(rayWeight, f_sub(d, dOld)), mass);
                    }
                    wOld = w; dOld = d;
                }
                const bool keep = live && ((int)kk >= firstIn - 1);
                res = keep ? res : 0.0f;
                rs = keep ? rs : inf;
                const unsigned rsBits = __float_as_uint(rs);
                const float m = __any_sync(0xffffffffu, (int)rsBits < 0) ? warp_min(rs) : __uint_as_float(__reduce_min_sync(0xffffffffu, rsBits));
                const bool anyDose = __any_sync(0xffffffffu, res > 0.0f);
                if (anyDose || q.writeAll) { iddP[off] = res; rsigP[off] = rs; }
                if (lane == 0) rowMinP[rmOff] = anyDose ? m : -m;
                off += (unsigned)R; rmOff += rmStride;
            }
        }
        wPrev = wLast;
        k += (unsigned)nb;
    }
#else
    // Four statically named slots (step s lives in slot s & 3, the loop is unrolled by four so that no register rotates at the back
    // edge): at step k the WEPL / density of step k + 3 are requested, the two LUT fetches (IDD at its WEPL, 1/X0 at its density) and
    // the scattering power of step k + 2 are issued, the step table entry of step k + 1 is loaded. Round 2's profile of the two-slot
    // version showed half of all stall samples on the first use of the 1/X0 fetch and of the step table: one step (~400 issue
    // slots shared by the scheduler's warps) does not cover a texture round trip.
    struct Slot { float w, rho, d, x, t; };
    Slot S0 = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f}, S1 = S0, S2 = S0, S3 = S0;
    float2 stNext = make_float2(0.0f, 0.0f);
    if (live && first < last) {
        S0.w = weplP[0]; S0.rho = densP[0];
        if (first + 1 < last) { S1.w = weplP[R]; S1.rho = densP[R]; }
        if (first + 2 < last) { S2.w = weplP[2 * R]; S2.rho = densP[2 * R]; }
        S0.d = tex2D<float>(iddTex, f_fma(S0.w, lp.energyScale, 0.5f), eRow);
        S0.x = tex1D<float>(rrlTex, f_fma(S0.rho, q.rrlScale, 0.5f));
        S0.t = scatPower(0.0f, S0.w);
        if (first + 1 < last) {
            S1.d = tex2D<float>(iddTex, f_fma(S1.w, lp.energyScale, 0.5f), eRow);
            S1.x = tex1D<float>(rrlTex, f_fma(S1.rho, q.rrlScale, 0.5f));
            S1.t = scatPower(S0.w, S1.w);
        }
        stNext = __ldg(stepTab);
    }
    // one step: C = slot of step k, N1 / N2 / N3 = slots of steps k + 1 / k + 2 / k + 3 (N3 is the slot step k - 1 has left)
    auto step = [&](const unsigned k, const Slot& C, const Slot& N1, Slot& N2, Slot& N3) {
        const float w = C.w, rho = C.rho, xC = C.x, tC = C.t, dC = C.d;
        const float2 st = stNext;                                  // 0.5 (vwx + vwy), stepVol of this step
        if (live && k + 1 < last) stNext = __ldg(stepTab + (k + 1 - first));
        if (live && k + 2 < last) {
            N2.d = tex2D<float>(iddTex, f_fma(N2.w, lp.energyScale, 0.5f), eRow);
            N2.x = tex1D<float>(rrlTex, f_fma(N2.rho, q.rrlScale, 0.5f));
            N2.t = scatPower(N1.w, N2.w);          // wOld of step k + 2 is the WEPL of step k + 1 (if the ray is still live there)
            if (k + 3 < last) { N3.w = weplP[off + 3 * R]; N3.rho = densP[off + 3 * R]; }
        }
        if (live) {
            if (w < lp.peakDepth) {
                const float rRl = f_mul(rho, xC);
                const float t1 = tC;
                const float thetaSq = f_mul(rRl, t1);
                sigmaSq = f_add(sigmaSq, f_add(incScat, incDiv));
                incinc = f_fma(stepLength, f_mul(stepLength, f_fma(rRl, t1, thetaSq)), incinc);
                incScat = f_add(incScat, incinc);
                incDiv = f_fma(airQuad, 2.0f, incDiv);
            } else if (q.nuclear != 3) {       // the empirical fix of the sigma dip after the peak is compiled out for GAUSS_FIT (cu:300-302)
                sigmaSq = f_fma(f_mul(f_add(incScat, incDiv), -1.5f), rho, sigmaSq);
            }
            rs = f_div(st.x, f_mul(f_add(f_sqrt(sigmaSq), q.sigmaDelta), 1.41421356f));
            const bool within = !(w > thr);
            live = within && (k != afterLast);
            afterLast = within ? afterLast : k;
            const float vol = st.y;
            const float mass = q.doseToMedium ? f_mul(rho, vol) : f_mul(f_sub(w, wOld), vol);
            const float d = dC;
            if (mass > 1e-2f) {
                if (q.nuclear) {
                    // the share of the step's dose the model attributes to the nuclear halo is taken out of the primary dose (cu:320-330);
                    // the halo itself never reaches the dose grid as the reference is shipped (nucMemoryStep = 0, rtdose.h)
                    const float nw = tex2D<float>(nucTex, f_fma(f_mul(f_add(w, wOld), 0.5f), lp.energyScale, 0.5f), eRow);
                    res = f_div(f_mul(f_mul(f_sub(1.0f, nw), rayWeight), f_sub(d, dOld)), mass);
                } else {
                    res = f_div(f_mul(rayWeight, f_sub(d, dOld)), mass);
                }
            }
            wOld = w; dOld = d;
        }
        const bool keep = live && ((int)k >= firstIn - 1);
        res = keep ? res : 0.0f;
        rs = keep ? rs : inf;
        // row minimum of rSigmaEff for the tile radius; its sign carries "some voxel of this row has dose > 0"
        // (a tile without dose is skipped by the superposition exactly like the reference's __syncthreads_or, cuh:456)
        // rs is positive or +inf: its bit pattern orders like an unsigned integer -> one REDUX instead of 5 shuffle+min pairs
        const unsigned rsBits = __float_as_uint(rs);
        const float m = __any_sync(0xffffffffu, (int)rsBits < 0) ? warp_min(rs) : __uint_as_float(__reduce_min_sync(0xffffffffu, rsBits));
        const bool anyDose = __any_sync(0xffffffffu, res > 0.0f);
        // A row segment without dose contributes nothing to the superposition whatever its sigma, and the tile radius needs only
        // its minimum (rowMin): its 2 x 128 B are not written. If another row of the same 32 x 8 tile has dose at this step the
        // superposition does read the segment - tile_radius_kernel, which looks at the tile's eight row minima anyway, then
        // fills it with (0, +inf). On C3 this removes ~60 % of the kernel's HBM writes.
        if (anyDose || q.writeAll) { iddP[off] = res; rsigP[off] = rs; }
        if (lane == 0) rowMinP[rmOff] = anyDose ? m : -m;
        off += (unsigned)R; rmOff += rmStride;
    };
    unsigned k = first;
    // a ray never comes back to life: once no lane of the warp is live (rays outside the field; every ray past its range)
    // the rest of the row is "no dose, sigma undefined": only its row minimum is written
    for (;;) {
        if (!(k < last && __any_sync(0xffffffffu, live))) break;
        step(k, S0, S1, S2, S3); ++k;
        if (!(k < last && __any_sync(0xffffffffu, live))) break;
        step(k, S1, S2, S3, S0); ++k;
        if (!(k < last && __any_sync(0xffffffffu, live))) break;
        step(k, S2, S3, S0, S1); ++k;
        if (!(k < last && __any_sync(0xffffffffu, live))) break;
        step(k, S3, S0, S1, S2); ++k;
    }
#endif
    if (q.writeAll) {
        for (unsigned kk = k; kk < last; ++kk) { iddP[off + (kk - k) * (unsigned)R] = 0.0f; rsigP[off + (kk - k) * (unsigned)R] = inf; }
    }
    // (the row minima of the steps after the segment's last live one keep beam_clear_kernel's -inf)
    firstPassive[(size_t)l * R + ray] = (int)afterLast;
    const int wm = warp_max_i((int)afterLast);
    if (lane == 0) atomicMax(layerFirstPassive + l, wm);
  }
}

// ---------------------------------------------------------------------------------------------------------
// Tile radius (own radius of every live 32x8 tile) + per-layer 34-bin histogram. One thread per (tile, slice);
// grid = (ceil(S*tiles/256), L): sized for the worst case, the beam's slice count comes from BeamDev.
// Also completes the sparse output of idd_sigma_kernel: in a tile that has dose, the row segments that have none were not written
// and get (0, +inf) here.
__global__ void __launch_bounds__(256)
tile_radius_kernel(const float* __restrict__ rowMin, const int* __restrict__ layerFirstPassive, const BeamDev* __restrict__ bd,
                   const int Rx, const int Ry, const float ksCutoff, const int fillRows, uint8_t* __restrict__ radMap, int* __restrict__ hist,
                   float* __restrict__ idd, float* __restrict__ rsig) {
    __shared__ int sHist[34];
    const int nz = bd->nz, first = bd->bfi;
    const int tilesX = Rx / 32, tilesY = Ry / 8, tiles = tilesX * tilesY;
    if ((long long)blockIdx.x * 256 >= (long long)nz * tiles) return;        // block-uniform
    const int l = blockIdx.y;
    if (threadIdx.x < 34) sHist[threadIdx.x] = 0;
    __syncthreads();
    const int nLive = min(max(layerFirstPassive[l] - first, 0), nz);
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx < nz * tiles) {
        const int kk = idx / tiles, t = idx % tiles, tx = t % tilesX, ty = t / tilesX;
        uint8_t r8 = 255;
        if (kk < nLive) {
            const float* p = rowMin + (((size_t)l * nz + kk) * Ry + ty * 8) * tilesX + tx;
            float v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) v[i] = p[(size_t)i * tilesX];
            float m = fabsf(v[0]);
            bool anyDose = v[0] > 0.0f;
#pragma unroll
            for (int i = 1; i < 8; ++i) { anyDose |= v[i] > 0.0f; if (fabsf(v[i]) < m) m = fabsf(v[i]); }
            // kernel_wrapper.cuh:303
            const unsigned rad = min((unsigned)f_toint_rz(f_add(f_div(ksCutoff, f_mul(f_sqrt(2.0f), m)), 0.5f)), 33u);
            r8 = anyDose ? (uint8_t)rad : (uint8_t)254;   // 254: live tile (counted in the histogram) that deposits nothing
            atomicAdd(&sHist[rad], 1);
            if (anyDose && fillRows) {
                const float inf = __int_as_float(0x7f800000);
                const size_t base = ((size_t)l * nz + kk) * ((size_t)Rx * Ry) + (size_t)(ty * 8) * Rx + tx * 32;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    if (v[i] > 0.0f) continue;
                    float4* a = reinterpret_cast<float4*>(idd + base + (size_t)i * Rx);
                    float4* b = reinterpret_cast<float4*>(rsig + base + (size_t)i * Rx);
#pragma unroll
                    for (int c = 0; c < 8; ++c) { a[c] = make_float4(0.0f, 0.0f, 0.0f, 0.0f); b[c] = make_float4(inf, inf, inf, inf); }
                }
            }
        }
        radMap[((size_t)l * nz + kk) * tiles + t] = r8;
    }
    __syncthreads();
    if (threadIdx.x < 34 && sHist[threadIdx.x]) atomicAdd(hist + l * 34 + threadIdx.x, sHist[threadIdx.x]);
}

// Batching rule of the reference's host loop (kernel_wrapper.cu:966-976) evaluated on the device, plus the
// work-list layout: tiles are ordered by descending applied radius (heavy work first).
// block = 64 threads, thread l handles layer l. stats: [0]=error flag (radius>32), [1]=total live tiles.
__global__ void batch_kernel(const int* __restrict__ hist, const int L, const int minTiles, int* __restrict__ batchRad,
                             int* __restrict__ rbCount /*[33]*/, int* __restrict__ stats) {
    __shared__ int sCount[33];
    if (threadIdx.x < 33) sCount[threadIdx.x] = 0;
    __syncthreads();
    for (int l = threadIdx.x; l < L; l += blockDim.x) {
        const int* c = hist + l * 34;
        int* rb = batchRad + l * 34;
        int layerMax = 0;
        for (int i = 0; i < 34; ++i) { rb[i] = i; if (c[i] > 0) layerMax = i; }
        if (c[33] > 0) atomicExch(&stats[0], 1);
        int rec = min(layerMax, 32), acc = 0;
        atomicAdd(&sCount[0], c[0]);
        for (int rad = min(layerMax, 32); rad > 0; --rad) {
            acc += c[rad];
            rb[rad] = rec;
            atomicAdd(&sCount[rec], c[rad]);
            if (acc >= minTiles) { rec = rad - 1; acc = 0; }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        // offsets in descending-radius order: rbCount[r] becomes the start offset of radius r
        int off = 0;
        for (int r = 32; r >= 0; --r) { const int n = sCount[r]; rbCount[r] = off; off += n; }
        stats[1] = off;
    }
}

// Items (slice, tile) that have at least one layer with an applied radius above 16: the work list of ks2_kernel<32, .>,
// which runs only 3 warps per SM and must not spend them scanning the whole beam. Deepest slices first.
__global__ void __launch_bounds__(256)
big_items_kernel(const uint8_t* __restrict__ radMap, const int* __restrict__ batchRad, const int L, const BeamDev* __restrict__ bd, const int tiles,
                 const int* __restrict__ bigTiles /* batch_kernel's offset of radius 16 = tiles with a larger applied radius */,
                 int* __restrict__ count, uint32_t* __restrict__ items) {
    if (*bigTiles == 0) return;                       // shallow target: nothing to list
    const int nz = bd->nz;
    const int t = blockIdx.x * 256 + threadIdx.x;
    const int nItems = nz * tiles;
    if (t >= nItems) return;
    const int idx = nItems - 1 - t;
    bool big = false;
    for (int l = 0; l < L && !big; ++l) {
        const unsigned r = radMap[(size_t)l * nItems + idx];
        big = r <= 32u && batchRad[l * 34 + r] > 16;
    }
    if (big) items[atomicAdd(count, 1)] = (uint32_t)idx;
}


// ---------------------------------------------------------------------------------------------------------
// K4: BEV dose -> dose grid, gather with trilinear interpolation straight from the linear haloed BEV volume
// (no copy into a cudaArray, cu:1107-1141). The B200 texture unit's linear filter is reproduced exactly as fitted
// with tools/texprobe (profiles/r01_texture_filter_model.md): per axis q = floor((c-0.5)*256+0.5), texel q>>8,
// fraction q&255; eight INTEGER weights summing to 256 split z -> x -> y with round-half-up on the "+1" branch and
// round-half-down on the x0 branch's y split; texels outside [first, zEnd) x [0,H) x [0,W) read 0 (border mode).

__device__ __forceinline__ int tex_q(float c) {   // fixed-point texel coordinate, 8 fractional bits
    return __float2int_rd(fmaf(c - 0.5f, 256.0f, 0.5f));
}

__device__ __forceinline__ float bev_plane(const float* __restrict__ pl, const GridParams& g, int ix, int iy, int a, int b, int Wz) {
    // weights of the four texels of one z-plane (sum = Wz)
    const int x1 = (Wz * a + 128) >> 8, x0 = Wz - x1;
    const int w11 = (x1 * b + 128) >> 8, w01 = (x0 * b + 127) >> 8;
    const bool xa = (unsigned)ix < (unsigned)g.W, xb = (unsigned)(ix + 1) < (unsigned)g.W;
    const bool ya = (unsigned)iy < (unsigned)g.H, yb = (unsigned)(iy + 1) < (unsigned)g.H;
    const float* r0 = pl + (size_t)iy * g.W + ix;
    const float* r1 = r0 + g.W;
    const float v00 = (xa && ya) ? r0[0] : 0.0f, v10 = (xb && ya) ? r0[1] : 0.0f;
    const float v01 = (xa && yb) ? r1[0] : 0.0f, v11 = (xb && yb) ? r1[1] : 0.0f;
    return fmaf((float)(x0 - w01), v00, fmaf((float)(x1 - w11), v10, fmaf((float)w01, v01, (float)w11 * v11)));
}

// Per-beam fills, sized by the beam's own slice count: the slices [bfi, Zg) of the haloed BEV dose (the only ones the superposition
// writes and the transfer reads) = 0; the row minima of every (layer, slice, segment) = -inf ("no dose, sigma undefined": what a
// segment the depth scan never visits, or has left behind, means to tile_radius_kernel); firstPassive = 0 (dead rays).
__global__ void __launch_bounds__(256)
beam_clear_kernel(float* __restrict__ bev, const BeamDev* __restrict__ bd, const int planeFloats /* (Rx + 64)(Ry + 64), a multiple of 4 */,
                  float* __restrict__ rowMin, const int rowMinPerSlice /* L-independent: Ry * Rx / 32 */, const int L,
                  int* __restrict__ firstPassive, const int LR /* L * rays, a multiple of 4 */) {
    const int nz = bd->nz;
    if (nz <= 0) return;
    const size_t tid = (size_t)blockIdx.x * 256 + threadIdx.x, nth = (size_t)gridDim.x * 256;
    float4* p = reinterpret_cast<float4*>(bev + (size_t)bd->bfi * planeFloats);
    const size_t n4 = (size_t)nz * (planeFloats / 4);
    for (size_t i = tid; i < n4; i += nth) p[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    const float ninf = __int_as_float(0xff800000);
    float4* rm = reinterpret_cast<float4*>(rowMin);
    const size_t m4 = (size_t)L * nz * rowMinPerSlice / 4;       // rowMinPerSlice is a multiple of 8
    for (size_t i = tid; i < m4; i += nth) rm[i] = make_float4(ninf, ninf, ninf, ninf);
    int4* fp = reinterpret_cast<int4*>(firstPassive);
    for (size_t i = tid; i < (size_t)LR / 4; i += nth) fp[i] = make_int4(0, 0, 0, 0);
}

// Persistent: the grid is sized from the machine, the work units (32 x 8 dose columns x one chunk of the box's z range) from the
// transfer box in BeamDev, which only the device knows when the kernel is launched.
__global__ void __launch_bounds__(256)
bev_to_grid_kernel(const float* __restrict__ bev, float* __restrict__ dose, const BeamDev* __restrict__ bd, unsigned long long* __restrict__ modified) {
    __shared__ GridParams sg;
    if (threadIdx.x < sizeof(GridParams) / 4) reinterpret_cast<int*>(&sg)[threadIdx.x] = reinterpret_cast<const int*>(&bd->grid)[threadIdx.x];
    __syncthreads();
    const GridParams& g = sg;
    const int nUnits = g.tilesX * g.tilesY * g.zChunks;
    unsigned nMod = 0;
    const size_t zs = (size_t)g.nx * g.ny;
    const size_t plane = (size_t)g.W * g.H;
    const int D = g.zEnd - g.first;
    const float* base = bev + (size_t)g.first * plane;
    const int zPer = g.zChunks > 0 ? (g.maxZ - g.minZ + g.zChunks) / g.zChunks : 0;
    for (int unit = blockIdx.x; unit < nUnits; unit += gridDim.x) {
        const int bz = unit / (g.tilesX * g.tilesY), rem = unit - bz * (g.tilesX * g.tilesY), by = rem / g.tilesX, bx = rem - by * g.tilesX;
        const int x = g.minX + bx * 32 + (threadIdx.x & 31);
        const int y = g.minY + by * 8 + (threadIdx.x >> 5);
        if (y >= g.ny || x >= g.nx) continue;
        // init(i,j) (transfer_param_struct_div3.cu:22-25)
        const float fi = i2f(x), fj = i2f(y);
        const float sx = f_add(f_fma(g.t.coefIdxI.x, fi, f_mul(g.t.coefIdxJ.x, fj)), g.t.coefOffset.x);
        const float sy = f_add(f_fma(g.t.coefIdxI.y, fi, f_mul(g.t.coefIdxJ.y, fj)), g.t.coefOffset.y);
        const float sz = f_add(f_fma(g.t.coefIdxI.z, fi, f_mul(g.t.coefIdxJ.z, fj)), g.t.coefOffset.z);
        const int zBeg = g.minZ + bz * zPer, zEnd = min(zBeg + zPer - 1, g.maxZ);
        float* res = dose + ((size_t)zBeg * g.ny + y) * g.nx + x;
        for (int z = zBeg; z <= zEnd; ++z, res += zs) {
            // getFanIdx(k) (cu:27-34) + HALF
            const float fk = i2f(z);
            const float rx = f_fma(g.t.inc.x, fk, sx), ry = f_fma(g.t.inc.y, fk, sy), rz = f_fma(g.t.inc.z, fk, sz);
            const float mx = f_add(f_div(rz, f_sub(g.t.normDistX, rz)), 1.0f);
            const float my = f_add(f_div(rz, f_sub(g.t.normDistY, rz)), 1.0f);
            const float cx = f_add(f_fma(rx, mx, g.t.globalOffset.x), 0.5f);
            const float cy = f_add(f_fma(ry, my, g.t.globalOffset.y), 0.5f);
            const float cz = f_add(f_add(rz, g.t.globalOffset.z), 0.5f);
            // coarse reject (also keeps the fixed-point conversion in range)
            if (!(cx > -1.0f && cy > -1.0f && cz > -1.0f && cx < (float)(g.W + 1) && cy < (float)(g.H + 1) && cz < (float)(D + 1))) continue;
            const int qx = tex_q(cx), qy = tex_q(cy), qz = tex_q(cz);
            const int ix = qx >> 8, iy = qy >> 8, iz = qz >> 8;
            const int a = qx & 255, b = qy & 255, c = qz & 255;
            float acc = 0.0f;
            if ((unsigned)iz < (unsigned)D && c != 256) acc = bev_plane(base + (size_t)iz * plane, g, ix, iy, a, b, 256 - c);
            if ((unsigned)(iz + 1) < (unsigned)D && c != 0) acc += bev_plane(base + (size_t)(iz + 1) * plane, g, ix, iy, a, b, c);
            const float v = acc * (1.0f / 256.0f);
            if (v > 0.0f) { if (g.atomic) atomicAdd(res, v); else *res += v; ++nMod; }
        }
    }
    // voxels modified by this launch (the bytes a host-resident dose grid exchanges over PCIe): one atomic per warp
    if (modified) {
        nMod = __reduce_add_sync(0xffffffffu, nMod);
        if ((threadIdx.x & 31) == 0 && nMod) atomicAdd(modified, (unsigned long long)nMod);
    }
}

// Multi-GPU tail: dst[i] += src[i] where src[i] != 0; dst may be the caller's pinned host grid (mapped pointer), so only
// voxels that received dose cross PCIe. One warp-aggregated count of the voxels touched.
__global__ void __launch_bounds__(256)
add_nonzero_kernel(const float* __restrict__ src, float* __restrict__ dst, const size_t n, unsigned long long* __restrict__ modified) {
    unsigned nMod = 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float v = src[i];
        if (v != 0.0f) { dst[i] += v; ++nMod; }
    }
    nMod = __reduce_add_sync(0xffffffffu, nMod);
    if ((threadIdx.x & 31) == 0 && nMod && modified) atomicAdd(modified, (unsigned long long)nMod);
}

// Multi-GPU push (N > 1, --reduce push): a rank's partial dose slab is added onto the plan's ONE dose grid that lives on
// another GPU of the NVSwitch domain (peer mapping, rtd_shared_dose_open). Only float4 groups holding dose cross NVLink,
// each as one 16-byte red.global.add.v4.f32 (a warp's 32 groups are 512 contiguous bytes); several ranks may push at
// once. Needs src and dst 16-byte aligned; the host wrapper checks that and handles the (< 4 float) tail.
__global__ void __launch_bounds__(256)
push_nonzero_kernel(const float4* __restrict__ src, float4* __restrict__ dst, const size_t n4) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 v = __ldg(src + i);
        if (v.x != 0.0f || v.y != 0.0f || v.z != 0.0f || v.w != 0.0f) atomicAdd(dst + i, v);
    }
}
__global__ void push_nonzero_tail_kernel(const float* __restrict__ src, float* __restrict__ dst, const int n) {
    const int i = threadIdx.x;
    if (i < n && src[i] != 0.0f) atomicAdd(dst + i, src[i]);
}

// Stream-ordered barrier between the GPUs of a node (rtd_group_*): every rank's stream runs this one-thread kernel at the same
// point of its sequence. `ctr` lives behind the plan's dose grid on rank 0 (peer-mapped on the other ranks) and only ever counts
// up: the k-th barrier of the group is over when it has seen k * world arrivals. Everything a rank queued before the barrier
// (its zeroing of the grid, its reductions into the grid over NVLink) is ordered before its arrival (release at system scope), and
// nothing queued after it starts before every rank has arrived (acquire). Replaces two 4-byte NCCL all-reduces per plan - each a
// kernel launch plus a rendezvous - by one remote atomic and a few polls.
// The number of the barrier is kept in this rank's device memory (localEpoch) so that the launch takes the same arguments every
// time and a captured plan (CUDA graph) can contain it.
__global__ void group_barrier_kernel(unsigned long long* __restrict__ ctr, unsigned long long* __restrict__ localEpoch, const unsigned long long world,
                                     int* __restrict__ timedOut, const unsigned long long maxNs) {
    const unsigned long long epoch = *localEpoch + 1ull;
    *localEpoch = epoch;
    const unsigned long long target = epoch * world;
    __threadfence_system();
    asm volatile("red.release.sys.global.add.u64 [%0], 1;" :: "l"(ctr) : "memory");
    unsigned long long t0, t1, seen;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(ctr) : "memory");
        if (seen >= target) break;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > maxNs) { *timedOut = 1; break; }      // a rank never arrived (it failed earlier): give the GPU back
        __nanosleep(200);
    }
    __threadfence_system();
}

// ---------------------------------------------------------------------------------------------------------
// One-call host path, CT side (rtd_plan_dose with a page-locked CT): the GPU PULLS the texels a beam's tracer can touch
// straight out of the caller's host memory (mapped pointer, coalesced reads over PCIe) into the CT array (surface writes).
// What is pulled is a per-row x interval [lo, hi] for every (y, z) row of the CT - much closer to the oblique, divergent
// frustum than any set of axis-aligned boxes - and a row keeps the interval that is already resident from earlier beams.
//
// ct_mark_rows_kernel: the frustum's boundary surface is tiled by small patches (two consecutive steps x two adjacent
// rays on the perimeter of the ray grid; the first and last step for every 2 x 2 block of rays). Every sample position
// is trilinear-affine in (i, j, k), so a patch lies inside the bounding box of its 4 corner samples; every row that box
// touches gets the box's x extent. A row's interval between its boundary crossings then covers every interior sample of
// the row as well. Slack: texel pair of the linear filter (+1) and 2 texels for the tracer's fp32 position accumulation.
__device__ __forceinline__ void ct_sample_pos(const TraceParams& p, const float s1, const float s2, const int i, const int j, const int k, float (&u)[3]) {
    const float fi = (float)i, fj = (float)j, fk = (float)k;
    const float I[3] = {p.coefIdxI.x, p.coefIdxI.y, p.coefIdxI.z}, J[3] = {p.coefIdxJ.x, p.coefIdxJ.y, p.coefIdxJ.z};
    const float T[3] = {p.transl.x, p.transl.y, p.transl.z}, O[3] = {p.coefOffset.x, p.coefOffset.y, p.coefOffset.z};
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const float st = T[d] + I[d] * fi * s1 + J[d] * fj * s2;
        const float in = (O[d] - I[d] * fi / p.distX - J[d] * fj / p.distY) * p.delta.z;
        u[d] = st + fk * in;          // texture coordinate - 1/2: texels floor(u), floor(u) + 1
    }
}

__global__ void __launch_bounds__(256)
ct_mark_rows_kernel(const TraceParams p, const int Rx, const int Ry, const int S, const int nx, const int ny, const int nz,
                    int* __restrict__ needLo, int* __restrict__ needHi) {
    const int nPerim = 2 * (Rx + Ry);                 // perimeter ray pairs (r, next r), closed loop incl. corners twice: harmless
    const long long nSide = (long long)nPerim * max(S - 1, 1);
    const long long nFace = 2LL * max(Rx - 1, 1) * max(Ry - 1, 1);
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    if (idx >= nSide + nFace) return;
    int i0, j0, i1, j1, k0, k1;
    if (idx < nSide) {
        const int e = (int)(idx % nPerim), k = (int)(idx / nPerim);
        k0 = k; k1 = min(k + 1, S - 1);
        if (e < Rx) { i0 = e; i1 = min(e + 1, Rx - 1); j0 = j1 = 0; }
        else if (e < 2 * Rx) { i0 = e - Rx; i1 = min(i0 + 1, Rx - 1); j0 = j1 = Ry - 1; }
        else if (e < 2 * Rx + Ry) { j0 = e - 2 * Rx; j1 = min(j0 + 1, Ry - 1); i0 = i1 = 0; }
        else { j0 = e - 2 * Rx - Ry; j1 = min(j0 + 1, Ry - 1); i0 = i1 = Rx - 1; }
    } else {
        long long f = idx - nSide;
        const int w = max(Rx - 1, 1), h = max(Ry - 1, 1);
        const int side = (int)(f / ((long long)w * h)); f -= (long long)side * w * h;
        i0 = (int)(f % w); j0 = (int)(f / w); i1 = min(i0 + 1, Rx - 1); j1 = min(j0 + 1, Ry - 1);
        k0 = k1 = side ? S - 1 : 0;
    }
    const float s1 = 1.0f - p.corner.z / p.distX, s2 = 1.0f - p.corner.z / p.distY;
    float mn[3] = {3.0e38f, 3.0e38f, 3.0e38f}, mx[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        float u[3];
        ct_sample_pos(p, s1, s2, (c & 1) ? i1 : i0, (c & 2) ? j1 : j0, (c & 4) ? k1 : k0, u);
#pragma unroll
        for (int d = 0; d < 3; ++d) { mn[d] = fminf(mn[d], u[d]); mx[d] = fmaxf(mx[d], u[d]); }
    }
    if (!(fabsf(mn[0]) < 1.0e9f && fabsf(mx[0]) < 1.0e9f && fabsf(mn[1]) < 1.0e9f && fabsf(mx[1]) < 1.0e9f && fabsf(mn[2]) < 1.0e9f && fabsf(mx[2]) < 1.0e9f)) return;
    const int x0 = max((int)floorf(mn[0]) - 2, 0), x1 = min((int)floorf(mx[0]) + 3, nx - 1);
    const int y0 = max((int)floorf(mn[1]) - 2, 0), y1 = min((int)floorf(mx[1]) + 3, ny - 1);
    const int z0 = max((int)floorf(mn[2]) - 2, 0), z1 = min((int)floorf(mx[2]) + 3, nz - 1);
    // x0 > x1 (patch beside the volume) still has to mark its rows: the row may be entered further along, and the hull of
    // all patches of the row is what counts; clamp the interval to the volume's edge instead of dropping it
    const int lo = min(x0, nx - 1), hi = max(x1, 0);
    for (int z = z0; z <= z1; ++z)
        for (int y = y0; y <= y1; ++y) {
            const int r = z * ny + y;
            if (lo < needLo[r]) atomicMin(needLo + r, lo);
            if (hi > needHi[r]) atomicMax(needHi + r, hi);
        }
}

// One warp per CT row: copies the part of [needLo, needHi] that is not resident yet from the caller's (mapped) host CT
// into the array, 128 contiguous bytes per warp access, and extends the row's resident interval.
__global__ void __launch_bounds__(256)
ct_pull_rows_kernel(const float* __restrict__ hostCt, cudaSurfaceObject_t ctSurf, const int nx, const int ny, const int nz,
                    const int* __restrict__ needLo, const int* __restrict__ needHi, int* __restrict__ resLo, int* __restrict__ resHi,
                    unsigned long long* __restrict__ pulledBytes) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= ny * nz) return;
    const int nl = needLo[row], nh = needHi[row];
    if (nh < nl) return;
    const int rl = resLo[row], rh = resHi[row];
    const int y = row % ny, z = row / ny;
    const float* src = hostCt + (size_t)row * nx;
    unsigned n = 0;
    auto copy = [&](int a, int b) {                   // texels a..b, aligned down to 32 floats so that every access is one 128 B line
        for (int x = (a & ~31) + lane; x <= b; x += 8 * 32) {
            float v[8];                               // 8 independent 128 B reads in flight per warp (PCIe round trip ~2 us)
#pragma unroll
            for (int q = 0; q < 8; ++q) { const int xx = x + 32 * q; v[q] = (xx >= a && xx <= b) ? __ldcs(src + xx) : 0.0f; }
#pragma unroll
            for (int q = 0; q < 8; ++q) { const int xx = x + 32 * q; if (xx >= a && xx <= b) { surf3Dwrite(v[q], ctSurf, xx * 4, y, z); ++n; } }
        }
    };
    if (rh < rl) copy(nl, nh);
    else { if (nl < rl) copy(nl, rl - 1); if (nh > rh) copy(rh + 1, nh); }
    if (lane == 0 && (rh < rl || nl < rl || nh > rh)) { resLo[row] = (rh < rl) ? nl : min(nl, rl); resHi[row] = (rh < rl) ? nh : max(nh, rh); }
    n = __reduce_add_sync(0xffffffffu, n);
    if (lane == 0 && n) atomicAdd(pulledBytes, 4ull * n);
}

}  // namespace rtd
