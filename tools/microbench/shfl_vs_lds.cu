// Micro-benchmark behind DESIGN.md 4 ("Round 2 on the superposition"): what does it cost on sm_100a to move a lane's register-held
// output column (NR floats) to its two destination columns (a) by a 128-bit read-modify-write of a column-major shared-memory patch
// (what ks2_rmw does) or (b) by two warp shuffles per float into register-held columns of the receiving lanes, with and without a
// block of NF packed FMAs per step in front (the superposition's vertical pass)?
// Prints clocks per step for W warps per SM. Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o shfl_vs_lds shfl_vs_lds.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 d; asm("add.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

constexpr int PC = 44, COLS = 64;
template <int MODE, int NR, int NF>
__global__ void __launch_bounds__(NR > 24 ? 256 : 512) bench(float* out, long long* clocks, int iters, int rad) {
    extern __shared__ __align__(16) float smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* patch = smem + warp * (PC * COLS);
    for (int i = lane; i < PC * COLS; i += 32) patch[i] = 0.f;
    constexpr int NP = NR / 2;
    u64 acc[NP], own[NP], halo[NP], w[8];
#pragma unroll
    for (int m = 0; m < NP; ++m) { acc[m] = pk(lane * 0.001f + m, lane * 0.002f + 1.f + m); own[m] = 0ull; halo[m] = 0ull; }
#pragma unroll
    for (int k = 0; k < 8; ++k) w[k] = pk(1.0f + 1e-6f * (lane + k), 1.0f - 1e-6f * k);
    __syncthreads();
    const long long t0 = clock64();
    float* col0 = patch + (lane + 16) * PC;
    for (int it = 0; it < iters; ++it) {
#pragma unroll 1
        for (int j = 0; j <= rad; ++j) {
            if (NF) {
#pragma unroll
                for (int f = 0; f < NF; ++f) acc[f % NP] = fma2(acc[f % NP], w[f & 7], w[(f + 3) & 7]);
            }
            if (MODE == 0) {              // smem read-modify-write of columns x + j and x - j
#pragma unroll
                for (int s = 0; s < 2; ++s) {
                    float4* c = reinterpret_cast<float4*>(col0 + (s ? -j : j) * PC);
#pragma unroll
                    for (int q = 0; q < NP / 2; ++q) {
                        float4 v = c[q];
                        const u64 lo = add2(pk(v.x, v.y), acc[2 * q]), hi = add2(pk(v.z, v.w), acc[2 * q + 1]);
                        upk(lo, v.x, v.y); upk(hi, v.z, v.w);
                        c[q] = v;
                    }
                    __syncwarp();
                }
            } else {                      // shuffles into register-held columns
                const bool left = lane < 16;
                const int srcA = left ? lane + j : lane - j, srcB = (left ? lane - j : lane + j) & 31;
                const bool haloB = left ? (lane < j) : (lane + j >= 32);
                const u64 mI = haloB ? 0ull : pk(1.f, 1.f), mH = haloB ? pk(1.f, 1.f) : 0ull;
#pragma unroll
                for (int m = 0; m < NP; ++m) {
                    float a0, a1; upk(acc[m], a0, a1);
                    const float p0 = __shfl_sync(0xffffffffu, a0, srcA), p1 = __shfl_sync(0xffffffffu, a1, srcA);
                    const float q0 = __shfl_sync(0xffffffffu, a0, srcB), q1 = __shfl_sync(0xffffffffu, a1, srcB);
                    own[m] = add2(own[m], pk(p0, p1));
                    const u64 b = pk(q0, q1);
                    own[m] = fma2(b, mI, own[m]);
                    halo[m] = fma2(b, mH, halo[m]);
                }
            }
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int m = 0; m < NP; ++m) { float a, b; upk(add2(add2(own[m], halo[m]), acc[m]), a, b); s += a + b; }
    for (int i = lane; i < PC * COLS; i += 32) s += patch[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) clocks[blockIdx.x] = t1 - t0;
}

template <int MODE, int NR, int NF>
void run(const char* name, int warps, int rad) {
    const int blocks = 148, iters = 200;
    float* out; long long* clk;
    cudaMalloc(&out, blocks * 512 * 4); cudaMalloc(&clk, blocks * 8);
    const size_t sh = (size_t)warps * PC * COLS * 4;
    cudaFuncSetAttribute(bench<MODE, NR, NF>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh);
    bench<MODE, NR, NF><<<blocks, warps * 32, sh>>>(out, clk, iters, rad);
    bench<MODE, NR, NF><<<blocks, warps * 32, sh>>>(out, clk, iters, rad);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < blocks; ++i) mean += double(h[i]); mean /= blocks;
    const double perStepSM = mean / (double(iters) * (rad + 1));       // clocks for ONE step of ALL warps of the SM
    printf("{\"case\": \"%s\", \"NR\": %d, \"NF\": %d, \"warps_per_sm\": %d, \"clk_per_step_all_warps\": %.1f, \"clk_per_warp_step\": %.1f, \"err\": \"%s\"}\n",
           name, NR, NF, warps, perStepSM, perStepSM / warps, cudaGetErrorString(e));
    cudaFree(out); cudaFree(clk);
}

int main() {
    for (int warps : {4, 8, 16}) {
        if (warps <= 8) {
            run<0, 40, 0>("rmw_smem", warps, 16);    run<1, 40, 0>("shfl_regs", warps, 16);
            run<0, 40, 132>("fma+rmw_smem", warps, 16); run<1, 40, 132>("fma+shfl_regs", warps, 16);
        }
        run<0, 24, 0>("rmw_smem", warps, 8);      run<1, 24, 0>("shfl_regs", warps, 8);
        run<0, 24, 68>("fma+rmw_smem", warps, 8);  run<1, 24, 68>("fma+shfl_regs", warps, 8);
    }
    return 0;
}
