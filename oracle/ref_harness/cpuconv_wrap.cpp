// Test / baseline infrastructure only (never on the product path).
// C entry point around the reference's OWN CPU convolution stage, compiled in place from
// /root/reference/src/cpu_convolution_1d.cpp (xConvCpuSparse :91, yConvCpuSparse :173) into oracle/_ref/libref_cpuconv.so
// by the Makefile next to this file. It is the only CPU implementation of any stage of the dose path that the reference
// holds (not part of its build: src/CMakeLists.txt:8,34), and BASELINE.json asks for it to be timed on the box's host
// cores beside the GPU numbers ("Baseline B"). The problem it is given is stage K3a, spot weights -> ray
// weights (gpu_convolution_2d.cu:16-71): per layer an x pass [Sy][Sx] -> [Sy][Rx] and a y pass [Sy][Rx] -> [Ry][Rx].
//
// The reference functions call exit(1) when rad > inOutOffset and do not bound-check their writes; both conditions are
// tested here first and reported as an error code instead.
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "cpu_convolution_1d.h"

extern "C" {

// weights [L][Sy][Sx]; rSig*[L] = 1 / (sqrt(2) sigma_eff), rad*[L]; spot (sx, sy) sits on ray (sx * dX + offX, sy * dY + offY);
// out [L][Ry][Rx] is overwritten. threads <= 0: all. Returns 0, or 1 + layer when that layer's taps would leave the ray grid.
int ref_cpuconv_layers(const float* weights, int L, int Sx, int Sy, const float* rSigX, const float* rSigY, const int* radX, const int* radY,
                       int Rx, int Ry, int offX, int offY, int dX, int dY, float* out, int threads) {
    for (int l = 0; l < L; ++l) {
        if (radX[l] < 0 || radY[l] < 0 || radX[l] > offX || radY[l] > offY) return 1 + l;
        if ((Sx - 1) * dX + offX + radX[l] >= Rx || (Sy - 1) * dY + offY + radY[l] >= Ry) return 1 + l;
    }
#ifdef _OPENMP
    const int nt = threads > 0 ? threads : omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(nt)
#endif
    for (int l = 0; l < L; ++l) {
        std::vector<float> interm((size_t)Sy * Rx, 0.0f);
        float* o = out + (size_t)l * Ry * Rx;
        std::memset(o, 0, sizeof(float) * (size_t)Ry * Rx);
        xConvCpuSparse(weights + (size_t)l * Sy * Sx, interm.data(), rSigX[l], (unsigned)radX[l], (unsigned)Sx, (unsigned)Rx, (unsigned)Sy, offX, dX);
        yConvCpuSparse(interm.data(), o, rSigY[l], (unsigned)radY[l], (unsigned)Sy, (unsigned)Rx, offY, dY);
    }
    return 0;
}

int ref_cpuconv_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

}  // extern "C"
