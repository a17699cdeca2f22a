// Issue rates on sm_100a of the instruction kinds the superposition is built from, alone and mixed (clocks per iteration per warp of a
// sub-partition with W warps per SM): FFMA, FFMA2 (fma.rn.f32x2), SHFL.IDX, LDS.32, LDS.128 + STS.128.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu ; results: profiles/r02_microbench_pipes.jsonl
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float d; asm volatile("fma.rn.ftz.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }

// MODE 0: 64 FFMA (16 chains)   1: 64 FFMA2 (16 chains)   2: 32 SHFL (independent)   3: 64 FFMA2 + 32 SHFL
//      4: 32 LDS.32   5: 64 FFMA2 + 32 LDS.32   6: 64 FFMA + 32 SHFL   7: 64 FFMA2 + 4 LDS.128 + 4 STS.128
template <int MODE>
__global__ void __launch_bounds__(512) bench(float* out, long long* clocks, int iters) {
    __shared__ __align__(16) float sm[16 * 512];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* my = sm + warp * 512;
    for (int i = lane; i < 512; i += 32) my[i] = i;
    float a[16]; u64 p[16]; float s[8];
#pragma unroll
    for (int k = 0; k < 16; ++k) { a[k] = lane + k; p[k] = pk(lane + k, lane - k); }
#pragma unroll
    for (int k = 0; k < 8; ++k) s[k] = lane * 3 + k;
    const float m0 = 1.0f + 1e-7f * lane, m1 = 1e-9f * lane; const u64 pm0 = pk(m0, m0), pm1 = pk(m1, m1);
    __syncthreads();
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
        if (MODE == 0 || MODE == 6) {
#pragma unroll
            for (int f = 0; f < 64; ++f) a[f & 15] = fma1(a[f & 15], m0, m1);
        }
        if (MODE == 1 || MODE == 3 || MODE == 5 || MODE == 7) {
#pragma unroll
            for (int f = 0; f < 64; ++f) p[f & 15] = fma2(p[f & 15], pm0, pm1);
        }
        if (MODE == 2 || MODE == 3 || MODE == 6) {
#pragma unroll
            for (int f = 0; f < 32; ++f) s[f & 7] = __shfl_sync(0xffffffffu, s[f & 7], (lane + it) & 31);
        }
        if (MODE == 4 || MODE == 5) {
#pragma unroll
            for (int f = 0; f < 32; ++f) { float v; asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(v) : "r"((unsigned)__cvta_generic_to_shared(my + ((f * 32 + lane + it) & 511)))); s[f & 7] += v; }
        }
        if (MODE == 7) {
            float4* c = reinterpret_cast<float4*>(my) + lane;
#pragma unroll
            for (int f = 0; f < 4; ++f) { float4 v = c[f * 32]; v.x += 1.f; c[f * 32] = v; }
        }
    }
    const long long t1 = clock64();
    float r = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) { float x, y; upk(p[k], x, y); r += a[k] + x + y; }
#pragma unroll
    for (int k = 0; k < 8; ++k) r += s[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r + my[lane];
    if (threadIdx.x == 0) clocks[blockIdx.x] = t1 - t0;
}
template <int MODE> void run(const char* name, int warps) {
    const int blocks = 148, iters = 2000;
    float* out; long long* clk;
    cudaMalloc(&out, blocks * 512 * 4); cudaMalloc(&clk, blocks * 8);
    bench<MODE><<<blocks, warps * 32>>>(out, clk, iters);
    bench<MODE><<<blocks, warps * 32>>>(out, clk, iters);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < blocks; ++i) mean += double(h[i]); mean /= blocks;
    printf("{\"case\": \"%s\", \"warps_per_sm\": %d, \"clk_per_iter_per_SM\": %.1f, \"clk_per_iter_per_warp_of_a_subpartition\": %.1f, \"err\": \"%s\"}\n", name, warps, mean / iters,
           mean / iters / (warps / 4.0), cudaGetErrorString(e));
    cudaFree(out); cudaFree(clk);
}
int main() {
    for (int w : {4, 8, 16}) {
        run<0>("64 FFMA", w); run<1>("64 FFMA2", w); run<2>("32 SHFL", w); run<3>("64 FFMA2 + 32 SHFL", w);
        run<4>("32 LDS.32", w); run<5>("64 FFMA2 + 32 LDS.32", w); run<6>("64 FFMA + 32 SHFL", w); run<7>("64 FFMA2 + 4 LDS.128 + 4 STS.128", w);
    }
    return 0;
}
