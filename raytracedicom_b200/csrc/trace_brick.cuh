// K1, variant B: the fan-beam density / WEPL tracer with the CT staged in shared memory by TMA instead of fetched through the texture
// pipe (north_star (1): "CT bricks staged by TMA or served from shared memory or texture").
//
// Reference: fillBevDensityAndSp (/root/reference/src/kernel_wrapper.cu:130-187) - one thread per ray, a serial loop over the depth
// steps with one tex3D (CT, linear filter) and two tex1D (HU -> density, HU -> stopping power) per step. Variant A (trace_bev_kernel,
// kernels.cuh) keeps the tex3D and is bound by the texture pipe (tex_throttle is its top stall, profiles/r01_ncu_full_rest.txt).
// Here:
//   * one CTA = a tile of 32 x 8 rays (a warp = 32 x-adjacent rays, so the density / WEPL rows it writes are whole 128-byte lines and
//     the LUT fetches of a warp stay coherent); thread = ray, marching its ray in depth like the reference's thread does, so the WEPL
//     sum has the reference's order by construction and needs no separate scan warp;
//   * the depth steps are cut into rounds of NS steps. Per round the CTA computes the lowest CT texel its 256 x NS sample positions
//     touch (the positions are monotonic in the step: a thread's extremes are its first and last sample of the round), and ONE thread
//     issues a cp.async.bulk.tensor.3d for the box that starts there from the fp32 CT in linear memory into shared memory,
//     completion on an mbarrier. TMA fills what lies outside the CT with zeros - the border addressing mode of the reference's
//     texture. The x origin is floored to a multiple of 4 texels: the innermost coordinate of a tensor copy must be 16-byte aligned
//     on this hardware (measured, tools/tmaprobe/README.md);
//   * the linear filter of the texture unit is applied by hand with the integer model fitted in round 1
//     (profiles/r01_texture_filter_model.md; same code shape as bev_plane in the transfer kernel): 8 integer weights that sum to 256,
//     split z -> x -> y. For INTEGER-valued texels below 2^15 (HU + 1000 of any DICOM CT with integer rescale) every product and
//     partial sum is an integer below 2^23, i.e. exact in fp32 in any order, and equals what the texture unit returns bit for bit.
//     The kernel checks every brick it loads; if it meets a texel that is not such an integer it raises flags[0] and the caller's
//     gated launch of variant A recomputes the beam through the texture unit (trace_bev_kernel(gate = flags)). flags[1] counts samples
//     that fell outside their brick (the host's box is meant to make that impossible); they are filtered from global memory.
//   * two CTAs per SM (brick <= ~100 KB): while one waits for its brick the other computes.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include "kernels.cuh"

namespace rtd {

struct BrickParams {
    int BX, BY, BZ;          // box of the tensor map (elements), BX a multiple of 4
    int NS;                  // depth steps per round
    int nx, ny, nz, pitch;   // CT dims and row pitch (floats) of the linear copy
    const float* ctLin;
};

constexpr int TRB_TX = 32, TRB_TY = 8, TRB_THREADS = TRB_TX * TRB_TY;

__device__ __forceinline__ uint32_t trb_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// One z-plane of the filter: four texels, weights sum to Wz (see bev_plane, kernels.cuh)
__device__ __forceinline__ float trb_plane(const float v00, const float v10, const float v01, const float v11, const int a, const int b, const int Wz) {
    const int x1 = (Wz * a + 128) >> 8, x0 = Wz - x1;
    const int w11 = (x1 * b + 128) >> 8, w01 = (x0 * b + 127) >> 8;
    return fmaf((float)(x0 - w01), v00, fmaf((float)(x1 - w11), v10, fmaf((float)w01, v01, (float)w11 * v11)));
}
__device__ __forceinline__ bool trb_is_hu_int(const float v) {      // integer-valued, 0 <= v < 2^15
    return truncf(v) == v && v >= 0.0f && v < 32768.0f;
}
__device__ __forceinline__ float trb_vox(const BrickParams& bp, const int x, const int y, const int z, bool& bad) {
    if ((unsigned)x >= (unsigned)bp.nx || (unsigned)y >= (unsigned)bp.ny || (unsigned)z >= (unsigned)bp.nz) return 0.0f;
    const float v = __ldg(bp.ctLin + ((size_t)z * bp.ny + y) * bp.pitch + x);
    bad |= !trb_is_hu_int(v);
    return v;
}

__global__ void __launch_bounds__(TRB_THREADS, 2)
trace_brick_kernel(const TraceParams p, const __grid_constant__ CUtensorMap tmap, const BrickParams bp,
                   cudaTextureObject_t densTex, cudaTextureObject_t spTex, cudaTextureObject_t densSpTex,
                   float* __restrict__ density, float* __restrict__ wepl, int* __restrict__ firstInside, int* __restrict__ firstOutside,
                   const int Rx, const int R, int* __restrict__ flags) {
    extern __shared__ __align__(128) unsigned char trbRaw[];
    __shared__ __align__(8) unsigned long long bar;
    __shared__ int sLo[2][TRB_TY][3];
    float* const brick = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(trbRaw) + 127) & ~uintptr_t(127));
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tilesX = Rx / TRB_TX;
    const int ix = (blockIdx.x % tilesX) * TRB_TX + lane, iy = (blockIdx.x / tilesX) * TRB_TY + warp;
    const int ray = iy * Rx + ix;
    const int S = (int)p.steps;
    const float fi = i2f(ix), fj = i2f(iy);

    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(trb_smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // getStart (density_and_sp_tracer_params.cu:15) + HALF, getInc (cu:17): the expressions of variant A, bit for bit
    const float s1 = f_sub(1.0f, f_div(p.corner.z, p.distX));
    const float s2 = f_sub(1.0f, f_div(p.corner.z, p.distY));
    float px = f_add(f_add(p.transl.x, f_fma(f_mul(p.coefIdxI.x, fi), s1, f_mul(f_mul(p.coefIdxJ.x, fj), s2))), 0.5f);
    float py = f_add(f_add(p.transl.y, f_fma(f_mul(p.coefIdxI.y, fi), s1, f_mul(f_mul(p.coefIdxJ.y, fj), s2))), 0.5f);
    float pz = f_add(f_add(f_fma(f_mul(p.coefIdxI.z, fi), s1, f_mul(f_mul(p.coefIdxJ.z, fj), s2)), p.transl.z), 0.5f);
    const float dx = f_mul(f_sub(f_sub(p.coefOffset.x, f_div(f_mul(p.coefIdxI.x, fi), p.distX)), f_div(f_mul(p.coefIdxJ.x, fj), p.distY)), p.delta.z);
    const float dy = f_mul(f_sub(f_sub(p.coefOffset.y, f_div(f_mul(p.coefIdxI.y, fi), p.distX)), f_div(f_mul(p.coefIdxJ.y, fj), p.distY)), p.delta.z);
    const float dz = f_mul(f_sub(f_sub(p.coefOffset.z, f_div(f_mul(p.coefIdxI.z, fi), p.distX)), f_div(f_mul(p.coefIdxJ.z, fj), p.distY)), p.delta.z);
    // stepLen (density_and_sp_tracer_params.cu:32-36)
    const float ddx = f_div(f_fma(p.delta.x, fi, p.corner.x), p.distX);
    const float ddy = f_div(f_fma(p.delta.y, fj, p.corner.y), p.distY);
    const float stepLen = f_mul(f_abs(p.delta.z), f_sqrt(f_fma(ddy, ddy, f_fma(ddx, ddx, 1.0f))));

    float cumSp = 0.0f, cumHu = 0.0f;
    int before = -1, last = -1;
    bool bad = false; int outside = 0;
    const int BX = bp.BX, BXY = bp.BX * bp.BY;
    const uint32_t brickBytes = (uint32_t)(BXY * bp.BZ) * 4u;
    const int brickVec = (BXY * bp.BZ) >> 2;
    size_t out = (size_t)ray;
    uint32_t phase = 0;
    __syncthreads();      // barrier initialised

    for (int k0 = 0, rnd = 0; k0 < S; k0 += bp.NS, ++rnd) {
        const int n = min(bp.NS, S - k0);
        {   // lowest texel index of this thread's samples of the round, per axis (the sample positions are monotonic in the step;
            // the last one is estimated, hence one texel of margin below - the other side is the host's business, BrickParams)
            const float m = i2f(n - 1);
            const int qx = tex_q(fminf(px, fmaf(dx, m, px))) >> 8, qy = tex_q(fminf(py, fmaf(dy, m, py))) >> 8, qz = tex_q(fminf(pz, fmaf(dz, m, pz))) >> 8;
            const int wx = __reduce_min_sync(0xffffffffu, qx), wy = __reduce_min_sync(0xffffffffu, qy), wz = __reduce_min_sync(0xffffffffu, qz);
            if (lane == 0) { sLo[rnd & 1][warp][0] = wx; sLo[rnd & 1][warp][1] = wy; sLo[rnd & 1][warp][2] = wz; }
        }
        __syncthreads();      // every thread has left the previous brick; sLo of this round is complete
        int ox = sLo[rnd & 1][0][0], oy = sLo[rnd & 1][0][1], oz = sLo[rnd & 1][0][2];
#pragma unroll
        for (int w = 1; w < TRB_TY; ++w) { ox = min(ox, sLo[rnd & 1][w][0]); oy = min(oy, sLo[rnd & 1][w][1]); oz = min(oz, sLo[rnd & 1][w][2]); }
        ox = (ox - 1) & ~3;   // 16-byte aligned innermost coordinate (floors for negative values too)
        oy -= 1; oz -= 1;
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(trb_smem_u32(&bar)), "r"(brickBytes) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                         :: "r"(trb_smem_u32(brick)), "l"(reinterpret_cast<uint64_t>(&tmap)), "r"(ox), "r"(oy), "r"(oz), "r"(trb_smem_u32(&bar)) : "memory");
        }
        {   // wait for the brick
            uint32_t done = 0;
            while (!done) {
                asm volatile("{\n\t.reg .pred q;\n\tmbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\tselp.u32 %0, 1, 0, q;\n\t}"
                             : "=r"(done) : "r"(trb_smem_u32(&bar)), "r"(phase) : "memory");
            }
            phase ^= 1u;
        }
        // every texel of the brick must be an integer below 2^15 for the hand-made filter to equal the texture unit's
        for (int i = threadIdx.x; i < brickVec; i += TRB_THREADS) {
            const float4 v = reinterpret_cast<const float4*>(brick)[i];
            bad |= !(trb_is_hu_int(v.x) && trb_is_hu_int(v.y) && trb_is_hu_int(v.z) && trb_is_hu_int(v.w));
        }
#pragma unroll 4
        for (int s = 0; s < n; ++s) {
            const int qx = tex_q(px), qy = tex_q(py), qz = tex_q(pz);
            const int a = qx & 255, b = qy & 255, c = qz & 255;
            const int gx = qx >> 8, gy = qy >> 8, gz = qz >> 8;
            const int lx = gx - ox, ly = gy - oy, lz = gz - oz;
            float acc;
            if ((unsigned)lx < (unsigned)(bp.BX - 1) && (unsigned)ly < (unsigned)(bp.BY - 1) && (unsigned)lz < (unsigned)(bp.BZ - 1)) {
                const float* t = brick + lz * BXY + ly * BX + lx;
                acc = trb_plane(t[0], t[1], t[BX], t[BX + 1], a, b, 256 - c);
                if (c != 0) acc += trb_plane(t[BXY], t[BXY + 1], t[BXY + BX], t[BXY + BX + 1], a, b, c);
            } else {
                ++outside;
                acc = trb_plane(trb_vox(bp, gx, gy, gz, bad), trb_vox(bp, gx + 1, gy, gz, bad), trb_vox(bp, gx, gy + 1, gz, bad),
                                trb_vox(bp, gx + 1, gy + 1, gz, bad), a, b, 256 - c);
                if (c != 0) acc += trb_plane(trb_vox(bp, gx, gy, gz + 1, bad), trb_vox(bp, gx + 1, gy, gz + 1, bad), trb_vox(bp, gx, gy + 1, gz + 1, bad),
                                             trb_vox(bp, gx + 1, gy + 1, gz + 1, bad), a, b, c);
            }
            const float hu = acc * (1.0f / 256.0f);
            float d, sp;
            if (densSpTex) {
                const float2 ds = tex1D<float2>(densSpTex, f_fma(hu, p.densityScale, 0.5f));
                d = ds.x; sp = ds.y;
            } else {
                d = tex1D<float>(densTex, f_fma(hu, p.densityScale, 0.5f));
                sp = tex1D<float>(spTex, f_fma(hu, p.spScale, 0.5f));
            }
            density[out] = d;
            cumHu = f_add(cumHu, hu);
            cumSp = f_fma(stepLen, sp, cumSp);
            if (cumHu < 150.0f) before = k0 + s;
            if (hu > 150.0f) last = k0 + s;
            wepl[out] = cumSp;
            out += (size_t)R;
            px = f_add(dx, px); py = f_add(dy, py); pz = f_add(dz, pz);
        }
    }
    firstInside[ray] = before + 1;
    firstOutside[ray] = last + 1;
    if (__any_sync(0xffffffffu, bad) && lane == 0) atomicOr(flags, 1);
    outside = __reduce_add_sync(0xffffffffu, outside);
    if (outside && lane == 0) atomicAdd(flags + 1, outside);
}

}  // namespace rtd
