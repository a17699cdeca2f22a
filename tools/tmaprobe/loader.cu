// Driver-API loader for the hand-written PTX variants of tools/tmaprobe/morph.py: loads <cubin>, encodes a 2-D fp32 tensor map here,
// launches k(tensorMap, out) with 128 threads and prints whether the 16 x 32 box at (0, 2) arrived.
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <vector>
#include <cuda.h>
int main(int argc, char** argv) {
    cuInit(0);
    CUdevice dev; cuDeviceGet(&dev, 0);
    CUcontext ctx; cuDevicePrimaryCtxRetain(&ctx, dev); cuCtxSetCurrent(ctx);
    FILE* f = fopen(argv[1], "rb"); if (!f) { printf("%s: no cubin\n", argv[1]); return 1; }
    std::vector<char> img(1 << 20); const size_t n = fread(img.data(), 1, img.size(), f); fclose(f); img.resize(n);
    CUmodule mod; CUresult r = cuModuleLoadData(&mod, img.data()); if (r) { printf("%s: load %d\n", argv[1], (int)r); return 1; }
    CUfunction fn; r = cuModuleGetFunction(&fn, mod, "k"); if (r) { printf("%s: getfn %d\n", argv[1], (int)r); return 1; }
    const int nx = 128, ny = 96;
    std::vector<float> h((size_t)nx * ny); for (size_t i = 0; i < h.size(); ++i) h[i] = float(i);
    CUdeviceptr d, o; cuMemAlloc(&d, h.size() * 4); cuMemcpyHtoD(d, h.data(), h.size() * 4); cuMemAlloc(&o, 512 * 4); cuMemsetD32(o, 0, 512);
    CUtensorMap tm; memset(&tm, 0, sizeof tm);
    const cuuint64_t dims[2] = {nx, ny}, strides[1] = {nx * 4};
    const cuuint32_t box[2] = {32, 16}, es[2] = {1, 1};
    r = cuTensorMapEncodeTiled(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    void* args[] = {&tm, &o};
    r = cuLaunchKernel(fn, 1, 1, 1, argc > 2 ? atoi(argv[2]) : 128, 1, 1, 8192, 0, args, nullptr);
    CUresult s = cuCtxSynchronize(); const char* es2 = nullptr; cuGetErrorString(s, &es2);
    std::vector<float> g(512, -2.f); cuMemcpyDtoH(g.data(), o, 512 * 4);
    printf("%-28s launch %d sync %d (%s) first %g %g %g (expect 256 257 258)\n", argv[1], (int)r, (int)s, es2 ? es2 : "", g[0], g[1], g[2]);
    return 0;
}
