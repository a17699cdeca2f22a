// What does the cold call pay for its workspaces? cudaMalloc / cudaFree of the same total (2.7 GB, C3's worst-case workspaces) as 75
// buffers of 36 MB, 6 of 450 MB, 1 of 2.7 GB; first touch excluded. Build: nvcc -O2 -o malloc_cost malloc_cost.cu
#include <chrono>
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
static double now() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main() {
    cudaFree(0);
    void* w; cudaMalloc(&w, 1 << 20); cudaFree(w);
    const size_t total = 2700ull << 20;
    for (int rep = 0; rep < 2; ++rep)
        for (int n : {75, 6, 1}) {
            std::vector<void*> p(n);
            const double t0 = now();
            for (int i = 0; i < n; ++i) if (cudaMalloc(&p[i], total / n) != cudaSuccess) { printf("malloc failed\n"); return 1; }
            const double t1 = now();
            for (int i = 0; i < n; ++i) cudaFree(p[i]);
            const double t2 = now();
            printf("{\"buffers\": %d, \"bytes_each\": %zu, \"malloc_ms\": %.3f, \"free_ms\": %.3f}\n", n, total / n, t1 - t0, t2 - t1);
        }
    return 0;
}
