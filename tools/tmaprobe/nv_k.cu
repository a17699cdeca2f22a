// The same kernel as morph.py's variants, written in CUDA C++ with inline PTX and compiled by nvcc: k(tensorMap, out).
#include <cuda.h>
#include <stdint.h>
#ifndef CX
#define CX 0
#endif
#ifndef NT
#define NT 128
#endif
#ifndef NB
#define NB 512
#endif
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
extern "C" __global__ void __launch_bounds__(NT) k(const __grid_constant__ CUtensorMap tmap, float* out) {
    __shared__ __align__(128) float brick[NB];
    __shared__ __align__(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(s32(&bar)));
#ifdef FENCE
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
#endif
    }
#ifdef PREWRITE
    for (int i = threadIdx.x; i < NB; i += NT) brick[i] = -1.f;
#endif
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(s32(&bar)), "r"(2048) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     :: "r"(s32(brick)), "l"(reinterpret_cast<uint64_t>(&tmap)), "r"(CX), "r"(2), "r"(s32(&bar)) : "memory");
    }
    uint32_t done = 0;
    while (!done) asm volatile("{\n\t.reg .pred q;\n\tmbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\tselp.u32 %0, 1, 0, q;\n\t}" : "=r"(done) : "r"(s32(&bar)), "r"(0u) : "memory");
    for (int i = threadIdx.x; i < 512; i += NT) out[i] = brick[i];
}
