// Stage-by-stage driver for the reference's OWN kernels and parameter classes, used to dump every
// intermediate of the dose path so the new engine and the CPU oracle can be checked stage by stage.
// TEST INFRASTRUCTURE ONLY (built into oracle/_ref/ref_stages). Nothing here is product code.
//
// It links the unmodified reference objects and launches, in the order of cudaWrapperProtons()
// (/root/reference/src/kernel_wrapper.cu:601-1218):
//   fillBevDensityAndSp (cu:130), sliceMinVar/sliceMaxVar (cuh:173-244), gpuConvolution2D (gpu_convolution_2d.cu:61),
//   fillIddAndSigma (cu:189), tileRadCalc (cuh:256), kernelSuperposition<r> (cuh:432), primTransfDiv (cu:69)
// with launch shapes and host glue restated from the cited lines, and writes raw little-endian dumps:
//
// usage: ref_stages <case.rtdcase> <outdir> [beamNo=0] [dump layers "all" | comma list]
//   <outdir>/meta.txt            key=value lines (dims, firstInside..., per-layer params, histograms, bbox)
//   density.f32 wepl.f32 [S][Ry][Rx]; first_inside.i32 first_outside.i32 [R]; wepl_min.f32 [S];
//   ray_weights.f32 [L][R]; idd_l<k>.f32 rsig_l<k>.f32 [S][R] (valid for steps [first, afterLast_l));
//   first_passive_l<k>.i32 [R]; bev_dose.f32 [Zg][Ry+64][Rx+64]; dose.f32 (beam's contribution on a zero grid)
#include <cstdio>
#include <cmath>
#include <iostream>
#include <fstream>
#include <set>
#include <sstream>
#include "helper_math.h"
#include "cuda_errchk.cuh"
#include "float3_from_fan_transform.cuh"
#include "float3_to_fan_transform.cuh"
#include "kernel_wrapper.cuh"
#include "vector_find.h"
#include "vector_interpolate.h"
#include "gpu_convolution_2d.cuh"
#include "case_io.h"

template <typename T>
static std::vector<T> d2h(const T* dev, size_t n) {
    std::vector<T> h(n);
    cudaErrchk(cudaMemcpy(h.data(), dev, n * sizeof(T), cudaMemcpyDeviceToHost));
    return h;
}

static cudaTextureObject_t makeTex(cudaArray* arr, cudaTextureAddressMode mode) {
    cudaResourceDesc r; memset(&r, 0, sizeof(r));
    r.resType = cudaResourceTypeArray; r.res.array.array = arr;
    cudaTextureDesc t; memset(&t, 0, sizeof(t));
    t.readMode = cudaReadModeElementType; t.normalizedCoords = false; t.filterMode = cudaFilterModeLinear;
    t.addressMode[0] = t.addressMode[1] = t.addressMode[2] = mode;
    cudaTextureObject_t o; cudaErrchk(cudaCreateTextureObject(&o, &r, &t, NULL));
    return o;
}

static cudaArray* make1D(const std::vector<float>& v) {
    cudaChannelFormatDesc d = cudaCreateChannelDesc<float>();
    cudaArray* a; cudaErrchk(cudaMallocArray(&a, &d, v.size()));
    cudaErrchk(cudaMemcpy2DToArray(a, 0, 0, v.data(), v.size() * 4, v.size() * 4, 1, cudaMemcpyHostToDevice));
    return a;
}

template <int R>
static void launchKs(int n, dim3 blk, float* idd, float* rs, float* out, int pitch, int2* idcs, int maxTiles, int* ctrs) {
    if (n > 0) kernelSuperposition<R><<<n, blk>>>(idd, rs, out, pitch, idcs, maxTiles, ctrs);
}
template <int R>
struct KsDispatch {
    static void run(const std::vector<int>& batched, dim3 blk, float* idd, float* rs, float* out, int pitch, int2* idcs, int maxTiles, int* ctrs) {
        launchKs<R>(batched[R], blk, idd, rs, out, pitch, idcs, maxTiles, ctrs);
        KsDispatch<R - 1>::run(batched, blk, idd, rs, out, pitch, idcs, maxTiles, ctrs);
    }
};
template <>
struct KsDispatch<-1> {
    static void run(const std::vector<int>&, dim3, float*, float*, float*, int, int2*, int, int*) {}
};

int main(int argc, char** argv) {
    if (argc < 3) { std::cerr << "usage: ref_stages case outdir [beam] [layers]\n"; return 2; }
    try {
        CaseFile c = readCase(argv[1]);
        const std::string out = std::string(argv[2]) + "/";
        const int beamNo = argc > 3 ? atoi(argv[3]) : 0;
        std::set<int> dumpLayers; bool dumpAll = true;
        if (argc > 4 && std::string(argv[4]) != "all") {
            dumpAll = false; std::stringstream ss(argv[4]); std::string tok;
            while (std::getline(ss, tok, ',')) dumpLayers.insert(atoi(tok.c_str()));
        }
        const CaseBeam& b = c.beams.at(beamNo);
        const CaseLuts& lut = c.luts;
        std::ofstream meta((out + "meta.txt").c_str());
        meta.precision(9);
        cudaErrchk(cudaFree(0));
        cudaChannelFormatDesc fd = cudaCreateChannelDesc<float>();

        // CT, LUT textures (cu:420-537)
        cudaArray* ctArr; cudaExtent ext = make_cudaExtent(c.ctDims[0], c.ctDims[1], c.ctDims[2]);
        cudaErrchk(cudaMalloc3DArray(&ctArr, &fd, ext));
        cudaMemcpy3DParms cp = {};
        cp.srcPtr = make_cudaPitchedPtr((void*)c.ct.data(), ext.width * 4, ext.width, ext.height);
        cp.dstArray = ctArr; cp.extent = ext; cp.kind = cudaMemcpyHostToDevice;
        cudaErrchk(cudaMemcpy3D(&cp));
        cudaTextureObject_t imTex = makeTex(ctArr, cudaAddressModeBorder);
        cudaArray* iddArr; cudaErrchk(cudaMallocArray(&iddArr, &fd, lut.nEnergySamples, lut.nEnergies));
        cudaErrchk(cudaMemcpy2DToArray(iddArr, 0, 0, lut.cidd.data(), lut.nEnergySamples * 4, lut.nEnergySamples * 4, lut.nEnergies, cudaMemcpyHostToDevice));
        cudaTextureObject_t iddTex = makeTex(iddArr, cudaAddressModeClamp);
        cudaTextureObject_t densTex = makeTex(make1D(lut.dens), cudaAddressModeClamp);
        cudaTextureObject_t spTex = makeTex(make1D(lut.sp), cudaAddressModeClamp);
        cudaTextureObject_t rrlTex = makeTex(make1D(lut.rrl), cudaAddressModeClamp);

        const size_t doseN = size_t(c.doseDims[0]) * c.doseDims[1] * c.doseDims[2];
        float* devDose; cudaErrchk(cudaMalloc(&devDose, doseN * 4)); cudaErrchk(cudaMemset(devDose, 0, doseN * 4));

        // Beam set-up (cu:612-663)
        const int L = b.L;
        std::vector<float> xs(L), ys(L);
        for (int i = 0; i < L; ++i) { xs[i] = b.sigmas[2 * i]; ys[i] = b.sigmas[2 * i + 1]; }
        const float2 maxSig = make_float2(findMax(xs), findMax(ys));
        const float3 spotDelta = make_float3(b.spotDelta[0], b.spotDelta[1], b.spotDelta[2]);
        const float3 spotOffset = make_float3(b.spotOffset[0], b.spotOffset[1], b.spotOffset[2]);
        const float2 sourceDist = make_float2(b.sourceDist[0], b.sourceDist[1]);
        const float3 rayRes = make_float3(b.raySpacing[0], b.raySpacing[1], spotDelta.z);
        const uint3 spotDims = make_uint3(b.Sx, b.Sy, b.L);
        int lSteps = int(ceil((spotOffset.x - (CONV_SIGMA_CUTOFF * maxSig.x + HALF * rayRes.x)) / rayRes.x));
        int bSteps = int(ceil((spotOffset.y - (CONV_SIGMA_CUTOFF * maxSig.y + HALF * rayRes.y)) / rayRes.y));
        int rSteps = int(floor(((spotDims.x - 1) * spotDelta.x + spotOffset.x + (CONV_SIGMA_CUTOFF * maxSig.x + HALF * rayRes.x)) / rayRes.x));
        int tSteps = int(floor(((spotDims.y - 1) * spotDelta.y + spotOffset.y + (CONV_SIGMA_CUTOFF * maxSig.y + HALF * rayRes.y)) / rayRes.y));
        float3 rayOffset = make_float3(rayRes.x * lSteps, rayRes.y * bSteps, spotOffset.z);
        Float3IdxTransform rayIdxToGantry(rayRes, rayOffset);
        Float3AffineTransform gtii(Matrix3x3(b.imM[0], b.imM[1], b.imM[2], b.imM[3], b.imM[4], b.imM[5], b.imM[6], b.imM[7], b.imM[8]), make_float3(b.imV[0], b.imV[1], b.imV[2]));
        Float3AffineTransform gtdi(Matrix3x3(b.doseM[0], b.doseM[1], b.doseM[2], b.doseM[3], b.doseM[4], b.doseM[5], b.doseM[6], b.doseM[7], b.doseM[8]), make_float3(b.doseV[0], b.doseV[1], b.doseV[2]));
        Float3FromFanTransform rayIdxToImIdx(rayIdxToGantry, sourceDist, gtii);
        const uint3 rayDims = make_uint3(roundTo(rSteps - lSteps + 1, superpTileX), roundTo(tSteps - bSteps + 1, superpTileY), L);
        const unsigned R = rayDims.x * rayDims.y, S = b.steps;
        meta << "lSteps=" << lSteps << "\nbSteps=" << bSteps << "\nrSteps=" << rSteps << "\ntSteps=" << tSteps
             << "\nRx=" << rayDims.x << "\nRy=" << rayDims.y << "\nS=" << S << "\nL=" << L << "\n";

        dim3 blk(32, 8), grid(rayDims.x / 32, rayDims.y / 8);
        float *dW, *dInterm, *dRayW, *dIdd, *dRs, *dDens, *dWepl, *dWeplMin; float2* dSig;
        int *dFi, *dFo, *dFp, *dBfi, *dBfo, *dLfp, *dCtr;
        cudaErrchk(cudaMalloc(&dW, spotDims.x * spotDims.y * L * 4));
        cudaErrchk(cudaMalloc(&dInterm, rayDims.x * spotDims.y * L * 4));
        cudaErrchk(cudaMalloc(&dRayW, size_t(R) * L * 4));
        cudaErrchk(cudaMalloc(&dSig, L * sizeof(float2)));
        cudaErrchk(cudaMalloc(&dIdd, size_t(R) * S * 4)); cudaErrchk(cudaMalloc(&dRs, size_t(R) * S * 4));
        cudaErrchk(cudaMalloc(&dDens, size_t(R) * S * 4)); cudaErrchk(cudaMalloc(&dWepl, size_t(R) * S * 4));
        cudaErrchk(cudaMemset(dIdd, 0, size_t(R) * S * 4)); cudaErrchk(cudaMemset(dRs, 0, size_t(R) * S * 4));
        cudaErrchk(cudaMalloc(&dFi, R * 4)); cudaErrchk(cudaMalloc(&dFo, R * 4)); cudaErrchk(cudaMalloc(&dFp, R * 4));
        cudaErrchk(cudaMalloc(&dBfi, 4)); cudaErrchk(cudaMalloc(&dBfo, 4)); cudaErrchk(cudaMalloc(&dLfp, 4));
        cudaErrchk(cudaMalloc(&dWeplMin, S * 4)); cudaErrchk(cudaMalloc(&dCtr, (maxSuperpR + 2) * 4));

        // K1 (cu:766-790)
        DensityAndSpTracerParams tp(lut.densScale, lut.spScale, S, rayIdxToImIdx);
        fillBevDensityAndSp<<<grid, blk>>>(dDens, dWepl, dFi, dFo, tp, imTex, densTex, spTex);
        cudaErrchk(cudaGetLastError());
        if (R < 1024) throw std::runtime_error("ray grid < 1024 rays: reference reductions read out of bounds (cuh:180,222)");
        sliceMinVar<int, 1024><<<1, 1024, 1024 * sizeof(int)>>>(dFi, dBfi, R);
        int bfi; cudaErrchk(cudaMemcpy(&bfi, dBfi, 4, cudaMemcpyDeviceToHost));
        float entryZ = float(bfi) * rayRes.z + rayOffset.z;
        sliceMaxVar<int, 1024><<<1, 1024, 1024 * sizeof(int)>>>(dFo, dBfo, R);
        int bfo; cudaErrchk(cudaMemcpy(&bfo, dBfo, 4, cudaMemcpyDeviceToHost));
        sliceMinVar<float, 128><<<dim3(1, 1, S), 128, 128 * sizeof(float)>>>(dWepl, dWeplMin, R);
        std::vector<float> weplMin = d2h(dWeplMin, S);
        writeRaw(out + "density.f32", d2h(dDens, size_t(R) * S).data(), size_t(R) * S * 4);
        writeRaw(out + "wepl.f32", d2h(dWepl, size_t(R) * S).data(), size_t(R) * S * 4);
        writeRaw(out + "first_inside.i32", d2h(dFi, R).data(), R * 4);
        writeRaw(out + "first_outside.i32", d2h(dFo, R).data(), R * 4);
        writeRaw(out + "wepl_min.f32", weplMin.data(), S * 4);

        // cu:792-849
        float maxEnergy = findMax<float>(b.energies);
        float maxEnergyIdx = findDecimalOrdered<float, float>(lut.energies, maxEnergy);
        float maxPeakDepth = vectorInterpolate<float, float>(lut.peakDepths, maxEnergyIdx);
        int firstPastCutoffAll = findFirstLargerOrdered<float>(weplMin, BP_DEPTH_CUTOFF * maxPeakDepth);
        const int Zg = min(firstPastCutoffAll, bfo);
        int fcp = 0;
        const int maxTiles = (Zg - bfi) * rayDims.x / superpTileX * rayDims.y / superpTileY;
        const uint3 bevDim = make_uint3(rayDims.x + 2 * maxSuperpR, rayDims.y + 2 * maxSuperpR, Zg);
        const size_t bevN = size_t(bevDim.x) * bevDim.y * bevDim.z;
        meta << "beamFirstInside=" << bfi << "\nbeamFirstOutside=" << bfo << "\nbeamFirstGuaranteedPassive=" << Zg << "\nentryZ=" << entryZ << "\n";
        if (Zg <= bfi) throw std::runtime_error("no live steps");
        float* dBev; cudaErrchk(cudaMalloc(&dBev, bevN * 4)); cudaErrchk(cudaMemset(dBev, 0, bevN * 4));
        int2* dIdcs; cudaErrchk(cudaMalloc(&dIdcs, size_t(maxSuperpR + 2) * maxTiles * sizeof(int2)));
        std::vector<float> eIdx(L), eScale(L), peak(L); std::vector<float2> entrySig(L);
        for (int l = 0; l < L; ++l) {
            eIdx[l] = findDecimalOrdered<float, float>(lut.energies, b.energies[l]);
            eScale[l] = vectorInterpolate<float, float>(lut.scaleFacts, eIdx[l]);
            peak[l] = vectorInterpolate<float, float>(lut.peakDepths, eIdx[l]);
            float2 co = FillIddAndSigmaParams::sigmaSqAirCoefs(peak[l]);
            float2 ss = make_float2(b.sigmas[2 * l], b.sigmas[2 * l + 1]);
            entrySig[l] = make_float2(sqrtf(co.x * entryZ * entryZ + co.y * entryZ + ss.x * ss.x),
                                      sqrtf(co.x * entryZ * entryZ + co.y * entryZ + ss.y * ss.y));
        }
        float2 pxSpMult = make_float2(1.0f - entryZ / sourceDist.x, 1.0f - entryZ / sourceDist.y);
        // K3a (cu:851-854)
        cudaErrchk(cudaMemcpy(dW, b.weights.data(), b.weights.size() * 4, cudaMemcpyHostToDevice));
        cudaErrchk(cudaMemcpy(dSig, entrySig.data(), L * sizeof(float2), cudaMemcpyHostToDevice));
        gpuConvolution2D(dW, dInterm, dRayW, dSig, spotDims, rayDims, spotDelta, spotOffset, rayRes, rayOffset, pxSpMult);
        cudaErrchk(cudaGetLastError());
        writeRaw(out + "ray_weights.f32", d2h(dRayW, size_t(R) * L).data(), size_t(R) * L * 4);

        // per-layer loop (cu:916-1056)
        long long updates = 0;
        for (int l = 0; l < L; ++l) {
            float spotDistInRays = spotDelta.x / b.raySpacing[0];
            unsigned localAfterLast = findFirstLargerOrdered<float>(weplMin, BP_DEPTH_CUTOFF * peak[l]);
            unsigned afterLast = min(localAfterLast, (unsigned)Zg);
            FillIddAndSigmaParams fp(eIdx[l], eScale[l], peak[l], entrySig[l].x * entrySig[l].x, lut.rrlScale, spotDistInRays, 0, bfi, afterLast, rayIdxToImIdx);
            fillIddAndSigma<<<grid, blk>>>(dDens, dWepl, dIdd, dRs, dRayW + size_t(l) * R, dFi, dFo, dFp, fp, iddTex, rrlTex);
            cudaErrchk(cudaGetLastError());
            sliceMaxVar<int, 1024><<<1, 1024, 1024 * sizeof(int)>>>(dFp, dLfp, R);
            int lfp; cudaErrchk(cudaMemcpy(&lfp, dLfp, 4, cudaMemcpyDeviceToHost));
            if (lfp > fcp) fcp = lfp;
            std::vector<int> ctr(maxSuperpR + 2, 0);
            cudaErrchk(cudaMemcpy(dCtr, ctr.data(), ctr.size() * 4, cudaMemcpyHostToDevice));
            int nLive = lfp - bfi;
            if (nLive > 0) {
                dim3 tg(rayDims.x / superpTileX, rayDims.y / superpTileY, nLive);
                tileRadCalc<4><<<tg, dim3(superpTileX, 4)>>>(dRs, bfi, dCtr, dIdcs, maxTiles);
                cudaErrchk(cudaGetLastError());
            }
            cudaErrchk(cudaMemcpy(ctr.data(), dCtr, ctr.size() * 4, cudaMemcpyDeviceToHost));
            updates += (long long)R * (nLive > 0 ? nLive : 0);
            meta << "layer" << l << "=energyIdx:" << eIdx[l] << ",scale:" << eScale[l] << ",peak:" << peak[l]
                 << ",entrySigX:" << entrySig[l].x << ",entrySigY:" << entrySig[l].y << ",afterLast:" << afterLast
                 << ",layerFirstPassive:" << lfp << "\nhist" << l << "=";
            for (size_t i = 0; i < ctr.size(); ++i) meta << ctr[i] << (i + 1 < ctr.size() ? "," : "\n");
            if (ctr[maxSuperpR + 1] > 0) throw std::runtime_error("radius > 32");
            if (dumpAll || dumpLayers.count(l)) {
                std::ostringstream n; n << l;
                writeRaw(out + "idd_l" + n.str() + ".f32", d2h(dIdd, size_t(R) * S).data(), size_t(R) * S * 4);
                writeRaw(out + "rsig_l" + n.str() + ".f32", d2h(dRs, size_t(R) * S).data(), size_t(R) * S * 4);
                writeRaw(out + "first_passive_l" + n.str() + ".i32", d2h(dFp, R).data(), R * 4);
            }
            int layerMax = 0;
            for (unsigned i = 0; i < maxSuperpR + 2; ++i) if (ctr[i] > 0) layerMax = i;
            int rec = layerMax;
            std::vector<int> batched(maxSuperpR + 1, 0);
            batched[0] = ctr[0];
            for (int rad = layerMax; rad > 0; --rad) {
                batched[rec] += ctr[rad];
                if (batched[rec] >= minTilesInBatch) rec = rad - 1;
            }
            KsDispatch<32>::run(batched, dim3(superpTileX, 8), dIdd, dRs, dBev, rayDims.x, dIdcs, maxTiles, dCtr);
            cudaErrchk(cudaGetLastError());
            cudaErrchk(cudaDeviceSynchronize());
        }
        meta << "beamFirstCalculatedPassive=" << fcp << "\nupdates=" << updates << "\n";
        writeRaw(out + "bev_dose.f32", d2h(dBev, bevN).data(), bevN * 4);

        // K4 (cu:1107-1218)
        cudaArray* bevArr; cudaExtent bext = make_cudaExtent(bevDim.x, bevDim.y, fcp - bfi);
        cudaErrchk(cudaMalloc3DArray(&bevArr, &fd, bext));
        cudaMemcpy3DParms bp = {};
        bp.srcPtr = make_cudaPitchedPtr((void*)(dBev + size_t(bfi) * bevDim.x * bevDim.y), bext.width * 4, bext.width, bext.height);
        bp.dstArray = bevArr; bp.extent = bext; bp.kind = cudaMemcpyDeviceToDevice;
        cudaErrchk(cudaMemcpy3D(&bp));
        cudaTextureObject_t bevTex = makeTex(bevArr, cudaAddressModeBorder);
        Float3FromFanTransform rayIdxToDoseIdx(rayIdxToGantry, sourceDist, gtdi);
        float3 mx = make_float3(-1.0f, -1.0f, -1.0f), mn = make_float3(100000.0f, 100000.0f, 100000.0f);
        float xv[2] = {-float(maxSuperpR), float(rayDims.x + maxSuperpR - 1)};
        float yv[2] = {-float(maxSuperpR), float(rayDims.y + maxSuperpR - 1)};
        float zv[2] = {float(bfi), float(fcp - 1)};
        for (int k = 0; k < 2; ++k) for (int j = 0; j < 2; ++j) for (int i = 0; i < 2; ++i) {
            float3 p = rayIdxToDoseIdx.transformPoint(make_float3(xv[i], yv[j], zv[k]));
            if (p.x > mx.x) mx.x = p.x; if (p.y > mx.y) mx.y = p.y; if (p.z > mx.z) mx.z = p.z;
            if (p.x < mn.x) mn.x = p.x; if (p.y < mn.y) mn.y = p.y; if (p.z < mn.z) mn.z = p.z;
        }
        int3 minIdx = make_int3(max(((int(floor(mn.x))) / 32) * 32, 0), max(int(floor(mn.y)), 0), max(int(floor(mn.z)), 0));
        int3 maxIdx = make_int3(min(int(ceil(mx.x)), c.doseDims[0] - 1), min(int(ceil(mx.y)), c.doseDims[1] - 1), min(int(ceil(mx.z)), c.doseDims[2] - 1));
        meta << "bboxMin=" << minIdx.x << "," << minIdx.y << "," << minIdx.z << "\nbboxMax=" << maxIdx.x << "," << maxIdx.y << "," << maxIdx.z << "\n";
        dim3 tgrid(roundTo(maxIdx.x - minIdx.x + 1, 32) / 32, roundTo(maxIdx.y - minIdx.y + 1, 8) / 8);
        TransferParamStructDiv3 ts(rayIdxToDoseIdx.invertAndShift(make_float3(float(maxSuperpR), float(maxSuperpR), -float(bfi))));
        primTransfDiv<<<tgrid, dim3(32, 8)>>>(devDose, ts, minIdx, maxIdx.z, make_uint3(c.doseDims[0], c.doseDims[1], c.doseDims[2]), bevTex);
        cudaErrchk(cudaGetLastError());
        cudaErrchk(cudaDeviceSynchronize());
        writeRaw(out + "dose.f32", d2h(devDose, doseN).data(), doseN * 4);
        meta.close();
        std::cout << "ref_stages ok: Rx=" << rayDims.x << " Ry=" << rayDims.y << " bfi=" << bfi << " Zg=" << Zg << " fcp=" << fcp << " updates=" << updates << std::endl;
    } catch (const std::exception& e) {
        std::cerr << "ref_stages failed: " << e.what() << std::endl; return 1;
    } catch (const char* e) {
        std::cerr << "ref_stages failed: " << e << std::endl; return 1;
    }
    return 0;
}
