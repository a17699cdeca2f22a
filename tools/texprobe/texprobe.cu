// Probe of the B200 texture unit's linear filter (test tooling, not product code): records what tex1D / tex3D return
// for known texels and coordinates so that the CPU oracle's software sampler (oracle/rtd_oracle.c, tex_frac) can be
// fitted to the hardware the reference's numerics depend on.
// usage: texprobe <outdir>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <string>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s line %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__global__ void probe1d(cudaTextureObject_t t, const float* x, float* out, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = tex1D<float>(t, x[i]);
}
__global__ void probe3d(cudaTextureObject_t t, const float* c, float* out, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = tex3D<float>(t, c[3 * i], c[3 * i + 1], c[3 * i + 2]);
}
__global__ void probe2d(cudaTextureObject_t t, const float* c, float* out, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = tex2D<float>(t, c[2 * i], c[2 * i + 1]);
}
static cudaTextureObject_t mk(cudaArray_t a, cudaTextureAddressMode m) {
    cudaResourceDesc r = {}; r.resType = cudaResourceTypeArray; r.res.array.array = a;
    cudaTextureDesc d = {}; d.readMode = cudaReadModeElementType; d.normalizedCoords = 0; d.filterMode = cudaFilterModeLinear;
    d.addressMode[0] = d.addressMode[1] = d.addressMode[2] = m;
    cudaTextureObject_t o; CK(cudaCreateTextureObject(&o, &r, &d, nullptr)); return o;
}
static void dump(const std::string& p, const void* d, size_t n) { FILE* f = fopen(p.c_str(), "wb"); fwrite(d, 1, n, f); fclose(f); }
static unsigned lcg(unsigned& s) { s = s * 1664525u + 1013904223u; return s; }
static float urand(unsigned& s) { return (lcg(s) >> 8) / 16777216.0f; }

int main(int argc, char** argv) {
    std::string out = argc > 1 ? argv[1] : ".";
    cudaChannelFormatDesc fd = cudaCreateChannelDesc<float>();
    // ---- 1-D ramp: tex(x) = floor + alpha
    const int N1 = 4096;
    std::vector<float> ramp(N1); for (int i = 0; i < N1; ++i) ramp[i] = (float)i;
    cudaArray_t a1; CK(cudaMallocArray(&a1, &fd, N1));
    CK(cudaMemcpy2DToArray(a1, 0, 0, ramp.data(), N1 * 4, N1 * 4, 1, cudaMemcpyHostToDevice));
    cudaTextureObject_t t1 = mk(a1, cudaAddressModeClamp);
    const float bases[6] = {0.0f, 1.0f, 10.0f, 100.0f, 1000.0f, 3000.0f};
    std::vector<float> xs;
    for (int b = 0; b < 6; ++b) for (int k = 0; k <= 65536 * 2; ++k) xs.push_back(bases[b] + (float)k / 65536.0f);
    unsigned s = 12345u;
    for (int k = 0; k < 1000000; ++k) xs.push_back(urand(s) * 4100.0f - 2.0f);
    int n = (int)xs.size();
    float *dx, *dy; CK(cudaMalloc(&dx, n * 4)); CK(cudaMalloc(&dy, n * 4));
    CK(cudaMemcpy(dx, xs.data(), n * 4, cudaMemcpyHostToDevice));
    probe1d<<<(n + 255) / 256, 256>>>(t1, dx, dy, n); CK(cudaDeviceSynchronize());
    std::vector<float> ys(n); CK(cudaMemcpy(ys.data(), dy, n * 4, cudaMemcpyDeviceToHost));
    dump(out + "/tp1d_x.f32", xs.data(), n * 4); dump(out + "/tp1d_y.f32", ys.data(), n * 4);
    // ---- 1-D random texels (lerp arithmetic)
    std::vector<float> rnd(N1); for (int i = 0; i < N1; ++i) rnd[i] = urand(s) * 3.0f;
    CK(cudaMemcpy2DToArray(a1, 0, 0, rnd.data(), N1 * 4, N1 * 4, 1, cudaMemcpyHostToDevice));
    probe1d<<<(n + 255) / 256, 256>>>(t1, dx, dy, n); CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(ys.data(), dy, n * 4, cudaMemcpyDeviceToHost));
    dump(out + "/tp1d_texels.f32", rnd.data(), N1 * 4); dump(out + "/tp1d_yr.f32", ys.data(), n * 4);
    // ---- 3-D random texels, border
    const int D = 16;
    std::vector<float> vol(D * D * D); for (auto& v : vol) v = urand(s) * 2000.0f;
    cudaArray_t a3; cudaExtent ext = make_cudaExtent(D, D, D); CK(cudaMalloc3DArray(&a3, &fd, ext));
    cudaMemcpy3DParms cp = {}; cp.srcPtr = make_cudaPitchedPtr(vol.data(), D * 4, D, D); cp.dstArray = a3; cp.extent = ext; cp.kind = cudaMemcpyHostToDevice;
    CK(cudaMemcpy3D(&cp));
    cudaTextureObject_t t3 = mk(a3, cudaAddressModeBorder);
    const int n3 = 1000000;
    std::vector<float> c3(3 * n3); for (auto& v : c3) v = urand(s) * 19.0f - 1.5f;
    float *dc, *do3; CK(cudaMalloc(&dc, n3 * 12)); CK(cudaMalloc(&do3, n3 * 4));
    CK(cudaMemcpy(dc, c3.data(), n3 * 12, cudaMemcpyHostToDevice));
    probe3d<<<(n3 + 255) / 256, 256>>>(t3, dc, do3, n3); CK(cudaDeviceSynchronize());
    std::vector<float> o3(n3); CK(cudaMemcpy(o3.data(), do3, n3 * 4, cudaMemcpyDeviceToHost));
    dump(out + "/tp3d_texels.f32", vol.data(), vol.size() * 4); dump(out + "/tp3d_c.f32", c3.data(), n3 * 12); dump(out + "/tp3d_y.f32", o3.data(), n3 * 4);
    // ---- 2-D random texels, clamp (IDD-like: 64 x 16)
    const int W2 = 64, H2 = 16;
    std::vector<float> im(W2 * H2); for (auto& v : im) v = urand(s) * 100.0f;
    cudaArray_t a2; CK(cudaMallocArray(&a2, &fd, W2, H2));
    CK(cudaMemcpy2DToArray(a2, 0, 0, im.data(), W2 * 4, W2 * 4, H2, cudaMemcpyHostToDevice));
    cudaTextureObject_t t2 = mk(a2, cudaAddressModeClamp);
    std::vector<float> c2(2 * n3); for (int i = 0; i < n3; ++i) { c2[2 * i] = urand(s) * 68.0f - 2.0f; c2[2 * i + 1] = urand(s) * 20.0f - 2.0f; }
    CK(cudaMemcpy(dc, c2.data(), n3 * 8, cudaMemcpyHostToDevice));
    probe2d<<<(n3 + 255) / 256, 256>>>(t2, dc, do3, n3); CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(o3.data(), do3, n3 * 4, cudaMemcpyDeviceToHost));
    dump(out + "/tp2d_texels.f32", im.data(), im.size() * 4); dump(out + "/tp2d_c.f32", c2.data(), n3 * 8); dump(out + "/tp2d_y.f32", o3.data(), n3 * 4);
    printf("texprobe ok\n");
    return 0;
}
