// Texture-filter probe 2 (test tooling): 2-D and 3-D textures whose texels are ramps along one axis, so a fetch
// returns index+fraction of that axis directly. usage: texprobe2 <outdir>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <string>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s line %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)
__global__ void probe3d(cudaTextureObject_t t, const float* c, float* out, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = tex3D<float>(t, c[3 * i], c[3 * i + 1], c[3 * i + 2]);
}
__global__ void probe2d(cudaTextureObject_t t, const float* c, float* out, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = tex2D<float>(t, c[3 * i], c[3 * i + 1]);
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
    const int D = 32, n = 400000;
    unsigned s = 777u;
    std::vector<float> c(3 * n);
    for (int i = 0; i < n; ++i) {
        if (i < 100000) {   // fine sweep along one axis, other two fixed at awkward fractions
            int ax = i / 33334; float t = 3.0f + (float)(i % 33334) / 16384.0f;
            c[3 * i] = 5.3137f; c[3 * i + 1] = 7.7719f; c[3 * i + 2] = 9.1213f; c[3 * i + ax] = t;
        } else { c[3 * i] = urand(s) * 30.0f + 1.0f; c[3 * i + 1] = urand(s) * 30.0f + 1.0f; c[3 * i + 2] = urand(s) * 30.0f + 1.0f; }
    }
    float *dc, *dout; CK(cudaMalloc(&dc, n * 12)); CK(cudaMalloc(&dout, n * 4));
    CK(cudaMemcpy(dc, c.data(), n * 12, cudaMemcpyHostToDevice));
    dump(out + "/p2_c.f32", c.data(), n * 12);
    std::vector<float> o(n);
    cudaArray_t a3; cudaExtent ext = make_cudaExtent(D, D, D); CK(cudaMalloc3DArray(&a3, &fd, ext));
    cudaTextureObject_t t3 = mk(a3, cudaAddressModeBorder);
    std::vector<float> vol(D * D * D);
    for (int mode = 0; mode < 5; ++mode) {   // 0,1,2: ramp along x,y,z; 3: x*y*z product; 4: random
        unsigned s2 = 4242u;
        for (int z = 0; z < D; ++z) for (int y = 0; y < D; ++y) for (int x = 0; x < D; ++x) {
            float v = mode == 0 ? x : mode == 1 ? y : mode == 2 ? z : mode == 3 ? (float)(x * y * z) : urand(s2) * 2000.0f;
            vol[(z * D + y) * D + x] = v;
        }
        cudaMemcpy3DParms cp = {}; cp.srcPtr = make_cudaPitchedPtr(vol.data(), D * 4, D, D); cp.dstArray = a3; cp.extent = ext; cp.kind = cudaMemcpyHostToDevice;
        CK(cudaMemcpy3D(&cp));
        probe3d<<<(n + 255) / 256, 256>>>(t3, dc, dout, n); CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(o.data(), dout, n * 4, cudaMemcpyDeviceToHost));
        dump(out + "/p2_3d_m" + std::to_string(mode) + ".f32", o.data(), n * 4);
        if (mode == 4) dump(out + "/p2_3d_texels.f32", vol.data(), vol.size() * 4);
    }
    cudaArray_t a2; CK(cudaMallocArray(&a2, &fd, D, D));
    cudaTextureObject_t t2 = mk(a2, cudaAddressModeClamp);
    std::vector<float> im(D * D);
    for (int mode = 0; mode < 4; ++mode) {   // 0,1: ramp x,y; 2: x*y; 3: random
        unsigned s2 = 999u;
        for (int y = 0; y < D; ++y) for (int x = 0; x < D; ++x) im[y * D + x] = mode == 0 ? x : mode == 1 ? y : mode == 2 ? (float)(x * y) : urand(s2) * 100.0f;
        CK(cudaMemcpy2DToArray(a2, 0, 0, im.data(), D * 4, D * 4, D, cudaMemcpyHostToDevice));
        probe2d<<<(n + 255) / 256, 256>>>(t2, dc, dout, n); CK(cudaDeviceSynchronize());
        CK(cudaMemcpy(o.data(), dout, n * 4, cudaMemcpyDeviceToHost));
        dump(out + "/p2_2d_m" + std::to_string(mode) + ".f32", o.data(), n * 4);
        if (mode == 3) dump(out + "/p2_2d_texels.f32", im.data(), im.size() * 4);
    }
    printf("texprobe2 ok\n");
    return 0;
}
