// librtdose.so: C ABI (include/rtdose.h) + host orchestration of the dose path.
// Replaces cudaWrapperProtons() (/root/reference/src/kernel_wrapper.cu:381-1369) with a context that owns all
// device memory (no per-beam cudaMalloc/cudaFree after the first call), issues every kernel on one stream and
// synchronises with the host twice per PLAN (the reference blocks ~3 times per energy layer, cu:916-976).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <string>
#include <vector>
#include <algorithm>
#include <mutex>
#include "../../include/rtdose.h"
#include "kernels.cuh"
#include "ks2.cuh"
#include "trace_brick.cuh"

#ifndef RTD_TRACER_BRICK_DEFAULT
#define RTD_TRACER_BRICK_DEFAULT 0      // what rtd_opts.tracerBrick = -1 means
#endif
#include "sharding.hpp"

using namespace rtd;

namespace {

// Workspaces come out of a few large chunks per context instead of one cudaMalloc each: on this pool's B200 boxes a cudaMalloc costs
// ~0.30 ms and a cudaFree ~0.27 ms WHATEVER the size (tools/microbench/malloc_cost.cu: 2.7 GB as 75 buffers 23 + 20 ms, as one buffer
// 0.8 + 1.0 ms), and a three-beam plan has ~75 buffers - two thirds of the cold drop-in call (e2e_cold) were allocator calls.
// Bump allocation, nothing is returned before the context dies; a buffer that has to grow takes a new region 1.5 x as large (so what
// it leaves behind stays below twice its final size). Chunks double from 256 MB to 2 GB.
template <bool PINNED>
struct ArenaT {
    std::vector<void*> chunks; char* cur = nullptr; size_t left = 0, next = PINNED ? (size_t(4) << 20) : (size_t(256) << 20);
    cudaError_t alloc(size_t bytes, void** out) {
        bytes = (bytes + 255) / 256 * 256;
        if (bytes > left) {
            size_t sz = std::max(bytes, next);
            void* c = nullptr;
            cudaError_t e = PINNED ? cudaMallocHost(&c, sz) : cudaMalloc(&c, sz);
            if (e != cudaSuccess && sz > bytes) { cudaGetLastError(); sz = bytes; e = PINNED ? cudaMallocHost(&c, sz) : cudaMalloc(&c, sz); }
            if (e != cudaSuccess) return e;
            chunks.push_back(c); cur = (char*)c; left = sz;
            next = std::min(next * 2, PINNED ? (size_t(64) << 20) : (size_t(2048) << 20));
        }
        *out = cur; cur += bytes; left -= bytes;
        return cudaSuccess;
    }
    // one chunk for everything a plan is about to ask for (no error if it cannot be had: alloc() then falls back to its own chunks)
    void reserve(size_t bytes) {
        bytes = (bytes + 255) / 256 * 256;
        if (bytes <= left) return;
        void* c = nullptr;
        if ((PINNED ? cudaMallocHost(&c, bytes) : cudaMalloc(&c, bytes)) != cudaSuccess) { cudaGetLastError(); return; }
        chunks.push_back(c); cur = (char*)c; left = bytes;
    }
    void release() { for (void* c : chunks) { if (PINNED) cudaFreeHost(c); else cudaFree(c); } chunks.clear(); cur = nullptr; left = 0; }
};
typedef ArenaT<false> DevArena;
typedef ArenaT<true> PinArena;

struct DevBuf {
    void* p = nullptr; size_t cap = 0; DevArena* arena = nullptr;
    static long& reallocs() { static long n = 0; return n; }      // any growth invalidates captured launch sequences
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        ++reallocs();
        if (arena) {
            const size_t want = std::max(bytes + bytes / 8, cap + cap / 2);
            void* q = nullptr;
            const cudaError_t e = arena->alloc(want, &q);
            if (e != cudaSuccess) return e;
            p = q; cap = want; return cudaSuccess;
        }
        if (p) { cudaError_t e = cudaFree(p); if (e != cudaSuccess) return e; p = nullptr; cap = 0; }
        size_t want = bytes + bytes / 8;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) return e;
        cap = want; return cudaSuccess;
    }
    void release() { if (p && !arena) cudaFree(p); p = nullptr; cap = 0; }
    template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};
struct PinBuf {
    void* p = nullptr; size_t cap = 0; PinArena* arena = nullptr;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (arena) {
            const size_t want = std::max(bytes + bytes / 8 + 64, cap + cap / 2);
            void* q = nullptr;
            const cudaError_t e = arena->alloc(want, &q);
            if (e != cudaSuccess) return e;
            p = q; cap = want; return cudaSuccess;
        }
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMallocHost(&p, bytes + bytes / 8 + 64);
        if (e != cudaSuccess) return e;
        cap = bytes + bytes / 8 + 64; return cudaSuccess;
    }
    void release() { if (p && !arena) cudaFreeHost(p); p = nullptr; cap = 0; }
    template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct TexArr {
    cudaArray_t arr = nullptr; cudaTextureObject_t tex = 0; cudaSurfaceObject_t surf = 0;
    void release() { if (tex) cudaDestroyTextureObject(tex); if (surf) cudaDestroySurfaceObject(surf); if (arr) cudaFreeArray(arr); tex = 0; surf = 0; arr = nullptr; }
};

struct Box { int lo[3], hi[3]; };   // inclusive voxel box (x, y, z)
inline bool boxInside(const Box& a, const Box& b) {   // a inside b
    for (int d = 0; d < 3; ++d) if (a.lo[d] < b.lo[d] || a.hi[d] > b.hi[d]) return false;
    return true;
}
inline int64_t boxVoxels(const Box& b) { return int64_t(b.hi[0] - b.lo[0] + 1) * (b.hi[1] - b.lo[1] + 1) * (b.hi[2] - b.lo[2] + 1); }

// Host buffers of a one-call plan (rtd_plan_dose / rtd_dose): only the sub-volumes the plan touches are moved.
struct HostIo {
    const float* ct = nullptr; float* dose = nullptr; bool zeroCopy = false; int64_t h2d = 0, d2h = 0;
    const float* ctMapped = nullptr;      // device-side alias of a page-locked ct: the CT rows are pulled by a kernel (ct_pull_rows_kernel)
    cudaEvent_t ctReady = nullptr;        // group CT relay: the CT sub-volumes of every beam of the call are in the array once this event has passed
};

struct BeamWs {
    // geometry (host, before anything runs)
    RayGrid grid; int S = 0, L = 0, R = 0, sx = 0, sy = 0;
    int l0 = 0, Lc = 0;   // computed layer range [l0, l0+Lc) of the beam's L layers (rtd_beam.layerBegin/End)
    float dist[2]; Affine gtii, gtdi;
    std::vector<LayerIn> lin;             // [Lc]
    // results of the device-side set-up (beam_setup_kernel), on the host after the final sync of the call
    int bfi = 0, bfo = 0, Zg = 0, nz = 0; bool live = false;
    float entryZ = 0, maxPeakDepth = 0;
    BeamDev hostDev;
    bool trBrick = false, trStood = false; int trOutside = 0;     // tracer variant B launched / its result stood / samples outside their brick
    rtd_beam_report rep; std::vector<rtd_layer_report> lrep;
    // device
    DevBuf density, wepl, firstIn, firstOut, weplMin, spotW, sigmas, rayW, layerIn, layerDev, beamDev, stepTab, idd, rsig, rowMin,
           firstPassive, lfp, batchRad, rbOffset, radMap, list, bev, segList;
    PinBuf hStage;   // staging of the uploads: spot weights | LayerIn[Lc]
    PinBuf hOut;     // BeamDev | LayerParams[Lc] | entry sigmas[Lc] | control block | batchRad[Lc*34]
    // lfp[L] | hist[L*34] | bigCursor[33] | stats[4] | work[4] | ksItems[4] | segCount[L] | iddQueue[4] | tracerFlags[4]
    size_t ctlInts() const { return size_t(Lc) + size_t(Lc) * 34 + 33 + 4 + 4 + 4 + size_t(Lc) + 4 + 4; }
    size_t outCtlOff() const { return (sizeof(BeamDev) + size_t(Lc) * (sizeof(LayerParams) + sizeof(float2)) + 15) / 16 * 16; }
    void adopt(DevArena* da, PinArena* pa) {
        DevBuf* all[] = {&density, &wepl, &firstIn, &firstOut, &weplMin, &spotW, &sigmas, &rayW, &layerIn, &layerDev, &beamDev, &stepTab, &idd, &rsig,
                         &rowMin, &firstPassive, &lfp, &batchRad, &rbOffset, &radMap, &list, &bev, &segList};
        for (DevBuf* b : all) b->arena = da;
        hStage.arena = pa; hOut.arena = pa;
    }
    void release() {
        DevBuf* all[] = {&density, &wepl, &firstIn, &firstOut, &weplMin, &spotW, &sigmas, &rayW, &layerIn, &layerDev, &beamDev, &stepTab, &idd, &rsig,
                         &rowMin, &firstPassive, &lfp, &batchRad, &rbOffset, &radMap, &list, &bev, &segList};
        for (DevBuf* b : all) b->release();
        hStage.release(); hOut.release();
    }
};

struct CallState { int launches = 0, hostSyncs = 1; bool serial = false, prepared = false, graph = false; cudaEvent_t evXferEnd = nullptr; };

}  // namespace

struct rtd_ctx {
    int device = 0; cudaStream_t stream = nullptr;
    bool ownStream = true;                // false: the caller's stream (rtd_create_on_stream), never destroyed here
    cudaStream_t cstream = nullptr;       // copy stream of the one-call host path (CT / dose-in sub-volumes)
    bool ctPartial = false;               // the CT array holds only the sub-volumes of the last rtd_plan_dose
    std::vector<Box> ctBoxes, doseBoxes;  // sub-volumes moved by the current call
    DevBuf ctRows;                        // int [4][nz * ny]: needLo, needHi (per beam), resLo, resHi (per call) of the pulled CT rows
    DevBuf xferCount; PinBuf hXferCount;  // dose voxels modified by the transfers of the current call
    std::string err;
    bool haveLuts = false, haveCt = false, haveDose = false;
    // LUTs (host copies of the small vectors the host logic needs)
    std::vector<float> energies, peakDepths, scaleFacts;
    int nEnergySamples = 0, nEnergies = 0; float densScale = 0, spScale = 0, rrlScale = 0;
    TexArr iddTex, densTex, spTex, densSpTex, rrlTex, ctTex, nucTex;   // nucTex: nuclear-halo weight (rtd_luts.nucWeightMatrix), optional   // densSpTex: density + SP as one float2 LUT when they share length and scale
    int ctDims[3] = {0, 0, 0}, doseDims[3] = {0, 0, 0};
    DevBuf ctLin; int ctPitch = 0;        // the CT once more as linear fp32 [nz][ny][ctPitch] (row pitch a multiple of 4 floats): source of the tracer's TMA bricks
    bool brickAttrSet = false;            // trace_brick_kernel's shared-memory attribute (set at first use: loading the kernel is not free)
    bool ctLinValid = false;              // ctLin holds what the CT array holds
    DevBuf dose;
    DevArena devArena; PinArena pinArena;      // every workspace of the context (not the IPC-shared group buffers, not the CUDA arrays)
    std::vector<BeamWs*> ws;
    int nBeamsLast = 0;
    std::vector<cudaEvent_t> events;
    std::vector<cudaStream_t> bstreams;   // per-beam streams (concurrent beams)
    int smCount = 148;
    int ks2Blocks8 = 0, ks2Blocks16 = 0, ks2Blocks24 = 0, ks2Blocks32 = 0, iddBlocks = 0;
    rtd_opts optsLast;
    // the ranks of a node that compute one plan together (rtd_group_*): ONE dose grid + control block on rank 0, mapped over CUDA IPC
    int groupRank = -1, groupWorld = 0; size_t groupFloats = 0;
    float* groupGrid = nullptr;                    // rank 0: own allocation; other ranks: peer mapping
    unsigned long long* groupCtr = nullptr;        // arrival counter behind the grid (monotonic: barrier k waits for k * world arrivals)
    DevBuf groupEpochDev;                          // this rank's barrier count (device memory: the barrier launch is replayable)
    cudaGraphExec_t groupGraphExec = nullptr; uint64_t groupGraphKey = 0;
    PinBuf groupFlag;                              // [0] = a barrier timed out
    std::vector<rtd_beam> groupBeams;
    // CT relay of the group (rtd_group_ct_create / _open): every CT box of a plan crosses PCIe ONCE on the node, into the staging buffer of
    // one rank, and the ranks that need it copy it from there over NVLink
    void* ctRelay = nullptr; size_t ctRelayBytes = 0; std::vector<void*> ctRelayPeers; bool ctRelayOpen = false;
    // captured launch sequence of the last device-resident plan (doseGraph)
    cudaGraphExec_t graphExec = nullptr; uint64_t graphKey = 0; bool graphBroken = false; long lutGeneration = 0;
    CallState graphCs;
};

namespace {

constexpr int KS2_WARPS_8 = 15, KS2_WARPS_16 = 8, KS2_WARPS_24 = 5, KS2_WARPS_32 = 3;   // warps per CTA: as many as the per-warp shared memory allows on one SM
constexpr size_t KS2_SMEM_8 = size_t(KS2_WARPS_8) * ks2_warp_floats(8) * 4, KS2_SMEM_16 = size_t(KS2_WARPS_16) * ks2_warp_floats(16) * 4,
                 KS2_SMEM_24 = size_t(KS2_WARPS_24) * ks2_warp_floats(24) * 4, KS2_SMEM_32 = size_t(KS2_WARPS_32) * ks2_warp_floats(32) * 4;
static_assert(KS2_SMEM_8 <= 227 * 1024 && KS2_SMEM_16 <= 227 * 1024 && KS2_SMEM_24 <= 227 * 1024 && KS2_SMEM_32 <= 227 * 1024, "one CTA per SM");

rtd_status fail(rtd_ctx* c, rtd_status s, const std::string& msg) { if (c) c->err = msg; return s; }

#define RTD_CUDA(call)                                                                                           \
    do {                                                                                                         \
        cudaError_t e_ = (call);                                                                                 \
        if (e_ != cudaSuccess) {                                                                                 \
            char b_[512]; snprintf(b_, sizeof b_, "cudaAssert: %s %s, line: %d", cudaGetErrorString(e_), __FILE__, __LINE__); \
            return fail(ctx, RTD_ERR_CUDA, b_);                                                                  \
        }                                                                                                        \
    } while (0)

cudaError_t makeTexture(TexArr& t, cudaTextureAddressMode mode) {
    cudaResourceDesc r; memset(&r, 0, sizeof r);
    r.resType = cudaResourceTypeArray; r.res.array.array = t.arr;
    cudaTextureDesc d; memset(&d, 0, sizeof d);
    d.readMode = cudaReadModeElementType; d.normalizedCoords = 0; d.filterMode = cudaFilterModeLinear;
    d.addressMode[0] = d.addressMode[1] = d.addressMode[2] = mode;
    return cudaCreateTextureObject(&t.tex, &r, &d, nullptr);
}

cudaError_t upload1D(TexArr& t, const float* v, int n, cudaStream_t s) {
    t.release();
    cudaChannelFormatDesc fd = cudaCreateChannelDesc<float>();
    cudaError_t e = cudaMallocArray(&t.arr, &fd, n);
    if (e != cudaSuccess) return e;
    e = cudaMemcpy2DToArrayAsync(t.arr, 0, 0, v, size_t(n) * 4, size_t(n) * 4, 1, cudaMemcpyHostToDevice, s);
    if (e != cudaSuccess) return e;
    return makeTexture(t, cudaAddressModeClamp);
}

cudaEvent_t ev(rtd_ctx* ctx, size_t i) {
    while (ctx->events.size() <= i) { cudaEvent_t e; cudaEventCreate(&e); ctx->events.push_back(e); }
    return ctx->events[i];
}

rtd_status ensureCtArray(rtd_ctx* ctx, int nx, int ny, int nz) {
    if (!ctx->ctTex.arr || ctx->ctDims[0] != nx || ctx->ctDims[1] != ny || ctx->ctDims[2] != nz) {
        RTD_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->ctTex.release();
        ctx->haveCt = false;
        cudaChannelFormatDesc fd = cudaCreateChannelDesc<float>();
        RTD_CUDA(cudaMalloc3DArray(&ctx->ctTex.arr, &fd, make_cudaExtent(nx, ny, nz), cudaArraySurfaceLoadStore));
        RTD_CUDA(makeTexture(ctx->ctTex, cudaAddressModeBorder));
        cudaResourceDesc rd; memset(&rd, 0, sizeof rd);
        rd.resType = cudaResourceTypeArray; rd.res.array.array = ctx->ctTex.arr;
        RTD_CUDA(cudaCreateSurfaceObject(&ctx->ctTex.surf, &rd));
        ctx->ctDims[0] = nx; ctx->ctDims[1] = ny; ctx->ctDims[2] = nz;
        ctx->ctPitch = (nx + 3) & ~3; ctx->ctLinValid = false;
    }
    return RTD_OK;
}

rtd_status setCtCommon(rtd_ctx* ctx, const float* src, int nx, int ny, int nz, cudaMemcpyKind kind) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!src || nx <= 0 || ny <= 0 || nz <= 0) return fail(ctx, RTD_ERR_INVALID, "rtd_set_ct: null pointer or non-positive dims");
    RTD_CUDA(cudaSetDevice(ctx->device));
    rtd_status s = ensureCtArray(ctx, nx, ny, nz);
    if (s != RTD_OK) return s;
    cudaMemcpy3DParms cp; memset(&cp, 0, sizeof cp);
    cp.srcPtr = make_cudaPitchedPtr((void*)src, size_t(nx) * 4, nx, ny);
    cp.dstArray = ctx->ctTex.arr; cp.extent = make_cudaExtent(nx, ny, nz); cp.kind = kind;
    RTD_CUDA(cudaMemcpy3DAsync(&cp, ctx->stream));
    ctx->ctLinValid = false;          // the linear copy of the TMA-brick tracer is refreshed on demand (ensureCtLinear)
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->haveCt = true; ctx->ctPartial = false;
    return RTD_OK;
}

// ---- tracer variant B (trace_brick.cuh) ------------------------------------------------------------------------------------------
inline bool brickWanted(const rtd_opts& o) { return o.tracerBrick > 0 || (o.tracerBrick < 0 && RTD_TRACER_BRICK_DEFAULT); }

// The resident CT once more as linear fp32 (row pitch a multiple of 4 floats), copied device to device from the texture's array the
// first time a plan asks for the brick tracer after rtd_set_ct (4 B / voxel more; outside any captured sequence: beamPrepare).
rtd_status ensureCtLinear(rtd_ctx* ctx) {
    if (ctx->ctLinValid || !ctx->haveCt || ctx->ctPartial) return RTD_OK;
    const size_t nx = ctx->ctDims[0], ny = ctx->ctDims[1], nz = ctx->ctDims[2];
    if (ctx->ctLin.ensure(size_t(ctx->ctPitch) * ny * nz * 4) != cudaSuccess) { cudaGetLastError(); return RTD_OK; }   // no room: texture tracer
    cudaMemcpy3DParms cp; memset(&cp, 0, sizeof cp);
    cp.srcArray = ctx->ctTex.arr;
    cp.dstPtr = make_cudaPitchedPtr(ctx->ctLin.p, size_t(ctx->ctPitch) * 4, nx, ny);
    cp.extent = make_cudaExtent(nx, ny, nz); cp.kind = cudaMemcpyDeviceToDevice;
    RTD_CUDA(cudaMemcpy3DAsync(&cp, ctx->stream));
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->ctLinValid = true;
    return RTD_OK;
}

// Box of the tensor map and steps per round for one beam.
// pos(i, j, k) = start(i, j) + k inc(i, j) with start and inc affine in i and in j (density_and_sp_tracer_params.cu:15-17), so over a
// tile of 32 x 8 rays and NS steps the extent of pos along CT axis a is at most
//   31 |cI_a| (|s1| + kmax |dz| / |distX|) + 7 |cJ_a| (|s2| + kmax |dz| / |distY|) + (NS - 1) max|inc_a|
// (+ 4: one texel of margin below, the neighbour texel of the filter above, and one of slack on either side for the rounding of the
// positions, which the kernel accumulates step by step; + 3 in x for the origin floored to a multiple of 4). The largest NS of 64, 32,
// 16, 8 whose brick fits the budget (two CTAs per SM) is taken.
constexpr size_t TRB_SMEM_BUDGET = 100 * 1024;
bool brickParamsOfBeam(const TraceParams& p, int Rx, int Ry, int S, BrickParams& bp) {
    const float cI[3] = {p.coefIdxI.x, p.coefIdxI.y, p.coefIdxI.z}, cJ[3] = {p.coefIdxJ.x, p.coefIdxJ.y, p.coefIdxJ.z};
    const float cO[3] = {p.coefOffset.x, p.coefOffset.y, p.coefOffset.z};
    const double dz = std::fabs(double(p.delta.z)), ax = 1.0 / std::fabs(double(p.distX)), ay = 1.0 / std::fabs(double(p.distY));   // 1 / inf = 0
    const double s1 = std::fabs(1.0 - double(p.corner.z) / double(p.distX)), s2 = std::fabs(1.0 - double(p.corner.z) / double(p.distY));
    for (int ns = 64; ns >= 8; ns >>= 1) {
        int B[3];
        bool ok = true;
        for (int a = 0; a < 3; ++a) {
            const double inc = dz * (std::fabs(double(cO[a])) + std::fabs(double(cI[a])) * (Rx - 1) * ax + std::fabs(double(cJ[a])) * (Ry - 1) * ay);
            const double ext = (TRB_TX - 1) * std::fabs(double(cI[a])) * (s1 + S * dz * ax) + (TRB_TY - 1) * std::fabs(double(cJ[a])) * (s2 + S * dz * ay) + (ns - 1) * inc;
            if (!(ext < 250.0)) { ok = false; break; }
            B[a] = int(std::ceil(ext)) + 4;
        }
        if (!ok) continue;
        B[0] = (B[0] + 3 + 3) & ~3;
        if (B[0] > 256 || B[1] > 256 || B[2] > 256) continue;
        if (size_t(B[0]) * B[1] * B[2] * 4 + 128 > TRB_SMEM_BUDGET) continue;
        bp.BX = B[0]; bp.BY = B[1]; bp.BZ = B[2]; bp.NS = ns;
        return true;
    }
    return false;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
bool makeCtTensorMap(rtd_ctx* ctx, const BrickParams& bp, CUtensorMap* out) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                 CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn fn = nullptr; static bool tried = false;
    if (!tried) {
        tried = true;
        void* f = nullptr; cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) == cudaSuccess && qr == cudaDriverEntryPointSuccess) fn = (EncodeFn)f;
        cudaGetLastError();
    }
    if (!fn) return false;
    const cuuint64_t dims[3] = {cuuint64_t(ctx->ctDims[0]), cuuint64_t(ctx->ctDims[1]), cuuint64_t(ctx->ctDims[2])};
    const cuuint64_t strides[2] = {cuuint64_t(ctx->ctPitch) * 4, cuuint64_t(ctx->ctPitch) * 4 * cuuint64_t(ctx->ctDims[1])};
    const cuuint32_t box[3] = {cuuint32_t(bp.BX), cuuint32_t(bp.BY), cuuint32_t(bp.BZ)}, es[3] = {1, 1, 1};
    return fn(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, ctx->ctLin.p, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// ---- sub-volume transfers of the one-call host path -------------------------------------------------------
// b minus o as at most 5 boxes, cut along z, then y (whole rows stay whole), then x - but never into rows shorter than 64 voxels, which
// starve the copy engine: such a remainder is kept uncut. An intersection below a fifth of b is not worth the extra copies.
void subtractBox(const Box& b, const Box& o, std::vector<Box>& out) {
    Box i;
    for (int d = 0; d < 3; ++d) { i.lo[d] = std::max(b.lo[d], o.lo[d]); i.hi[d] = std::min(b.hi[d], o.hi[d]); if (i.lo[d] > i.hi[d]) { out.push_back(b); return; } }
    if (boxVoxels(i) * 5 < boxVoxels(b)) { out.push_back(b); return; }
    Box r = b;
    for (int d = 2; d >= 1; --d) {
        if (r.lo[d] < i.lo[d]) { Box p = r; p.hi[d] = i.lo[d] - 1; out.push_back(p); r.lo[d] = i.lo[d]; }
        if (r.hi[d] > i.hi[d]) { Box p = r; p.lo[d] = i.hi[d] + 1; out.push_back(p); r.hi[d] = i.hi[d]; }
    }
    const int left = i.lo[0] - r.lo[0], right = r.hi[0] - i.hi[0];
    if (left == 0 && right == 0) return;                                  // the rest of b lies inside o
    if ((left == 0 || left >= 64) && (right == 0 || right >= 64)) {
        if (left) { Box p = r; p.hi[0] = i.lo[0] - 1; out.push_back(p); }
        if (right) { Box p = r; p.lo[0] = i.hi[0] + 1; out.push_back(p); }
    } else out.push_back(r);
}

// Queues the parts of a box that no already queued box holds (beams overlap around the isocentre: on a three-field head plan a
// quarter, on a five-field pelvis plan half of the summed box volume is overlap).
void addBox(std::vector<Box>& v, Box b, const int dims[3]) {
    for (int d = 0; d < 3; ++d) { b.lo[d] = std::max(b.lo[d], 0); b.hi[d] = std::min(b.hi[d], dims[d] - 1); if (b.lo[d] > b.hi[d]) return; }
    // full rows when most of a row is needed anyway: one contiguous run per row / slice for the copy engine
    if (int64_t(b.hi[0] - b.lo[0] + 1) * 4 >= int64_t(dims[0]) * 3) { b.lo[0] = 0; b.hi[0] = dims[0] - 1; }
    std::vector<Box> pieces{b}, next;
    const size_t nOld = v.size();
    for (size_t k = 0; k < nOld && !pieces.empty(); ++k) {
        next.clear();
        for (size_t q = 0; q < pieces.size(); ++q) {
            if (next.size() + (pieces.size() - q) > 12) next.push_back(pieces[q]);      // enough fragments: the rest goes as it is
            else subtractBox(pieces[q], v[k], next);
        }
        pieces.swap(next);
    }
    for (const Box& p : pieces) v.push_back(p);
}

// CT boxes that enclose every texel the tracer of this beam can fetch (trace_bev_kernel: pos = start(i,j) + 1/2 + k inc(i,j),
// density_and_sp_tracer_params.cu:15-17; linear filtering reads floor(pos - 1/2) and the texel after it). start and inc
// are affine in i and in j, so pos is affine in each of i, j, k separately and its extremes over a box of (i, j, k) are
// attained at the 8 corners. The frustum is cut into depth segments so that oblique beams do not drag their whole
// axis-aligned bounding box along.
void ctBoxesOfBeam(const TraceParams& p, int Rx, int Ry, int S, const int dims[3], std::vector<Box>& out) {
    const double s1 = 1.0 - double(p.corner.z) / double(p.distX), s2 = 1.0 - double(p.corner.z) / double(p.distY);
    auto start = [&](int i, int j, double r[3]) {
        const double I[3] = {p.coefIdxI.x, p.coefIdxI.y, p.coefIdxI.z}, J[3] = {p.coefIdxJ.x, p.coefIdxJ.y, p.coefIdxJ.z}, T[3] = {p.transl.x, p.transl.y, p.transl.z};
        for (int d = 0; d < 3; ++d) r[d] = T[d] + I[d] * i * s1 + J[d] * j * s2;
    };
    auto inc = [&](int i, int j, double r[3]) {
        const double I[3] = {p.coefIdxI.x, p.coefIdxI.y, p.coefIdxI.z}, J[3] = {p.coefIdxJ.x, p.coefIdxJ.y, p.coefIdxJ.z}, O[3] = {p.coefOffset.x, p.coefOffset.y, p.coefOffset.z};
        for (int d = 0; d < 3; ++d) r[d] = (O[d] - I[d] * i / double(p.distX) - J[d] * j / double(p.distY)) * double(p.delta.z);
    };
    auto boxesFor = [&](int segLen, std::vector<Box>& dst) {
        for (int k0 = 0; k0 < S; k0 += segLen) {
            const int k1 = std::min(k0 + segLen, S - 1);   // segments share their end planes
            double mn[3] = {1e30, 1e30, 1e30}, mx[3] = {-1e30, -1e30, -1e30};
            for (int c = 0; c < 8; ++c) {
                const int i = (c & 1) ? Rx - 1 : 0, j = (c & 2) ? Ry - 1 : 0, k = (c & 4) ? k1 : k0;
                double st[3], in[3]; start(i, j, st); inc(i, j, in);
                for (int d = 0; d < 3; ++d) { const double v = st[d] + k * in[d]; mn[d] = std::min(mn[d], v); mx[d] = std::max(mx[d], v); }
            }
            Box b; bool finite = true;
            for (int d = 0; d < 3; ++d) {
                if (!(std::fabs(mn[d]) < 1e9 && std::fabs(mx[d]) < 1e9)) { finite = false; break; }
                b.lo[d] = int(std::floor(mn[d])) - 2; b.hi[d] = int(std::floor(mx[d])) + 3;   // texel pair + fp32 accumulation slack
            }
            if (!finite) { b.lo[0] = b.lo[1] = b.lo[2] = 0; b.hi[0] = dims[0] - 1; b.hi[1] = dims[1] - 1; b.hi[2] = dims[2] - 1; }
            addBox(dst, b, dims);
        }
    };
    // Second family: DISJOINT slabs along the CT axis the beam mostly travels along (depth segments overlap each other by
    // their slack on every side, which for oblique beams doubles the bytes). Per step k the samples of the whole ray grid lie
    // inside the box of the 4 corner rays (pos is bilinear in (i, j)); a slab [a0, a1] of texels along the dominant axis can
    // only be touched by steps whose box reaches it (2 / 3 texels of slack as above); the transverse extent of those steps is
    // bounded by the corner rays at the first and last such step (pos is linear in k for a fixed ray).
    auto slabBoxes = [&](int width, std::vector<Box>& dst) -> bool {
        if (S < 2) return false;
        std::vector<double> mn(size_t(S) * 3, 1e30), mx(size_t(S) * 3, -1e30);
        double st[4][3], in[4][3];
        for (int c = 0; c < 4; ++c) { start((c & 1) ? Rx - 1 : 0, (c & 2) ? Ry - 1 : 0, st[c]); inc((c & 1) ? Rx - 1 : 0, (c & 2) ? Ry - 1 : 0, in[c]); }
        for (int k = 0; k < S; ++k)
            for (int c = 0; c < 4; ++c)
                for (int d = 0; d < 3; ++d) {
                    const double v = st[c][d] + k * in[c][d];
                    if (!(std::fabs(v) < 1e9)) return false;
                    mn[size_t(k) * 3 + d] = std::min(mn[size_t(k) * 3 + d], v); mx[size_t(k) * 3 + d] = std::max(mx[size_t(k) * 3 + d], v);
                }
        int a = 0;
        for (int d = 1; d < 3; ++d) if (std::fabs(in[0][d]) > std::fabs(in[0][a])) a = d;
        double lo = 1e30, hi = -1e30;
        for (int k = 0; k < S; ++k) { lo = std::min(lo, mn[size_t(k) * 3 + a]); hi = std::max(hi, mx[size_t(k) * 3 + a]); }
        const int loA = std::max(int(std::floor(lo)) - 2, 0), hiA = std::min(int(std::floor(hi)) + 3, dims[a] - 1);
        for (int a0 = loA; a0 <= hiA; a0 += width) {
            const int a1 = std::min(a0 + width - 1, hiA);
            int k0 = -1, k1 = -1;
            for (int k = 0; k < S; ++k)
                if (int(std::floor(mx[size_t(k) * 3 + a])) + 3 >= a0 && int(std::floor(mn[size_t(k) * 3 + a])) - 2 <= a1) { if (k0 < 0) k0 = k; k1 = k; }
            if (k0 < 0) continue;
            Box b;
            for (int d = 0; d < 3; ++d) {
                b.lo[d] = int(std::floor(std::min(mn[size_t(k0) * 3 + d], mn[size_t(k1) * 3 + d]))) - 2;
                b.hi[d] = int(std::floor(std::max(mx[size_t(k0) * 3 + d], mx[size_t(k1) * 3 + d]))) + 3;
            }
            b.lo[a] = a0; b.hi[a] = a1;
            addBox(dst, b, dims);
        }
        return true;
    };
    // take the candidate that moves the fewest voxels: depth segments of several lengths (short ones follow an oblique beam
    // closely but pay the slack on every box, long ones drag the beam's bounding box along) or axis slabs (rows of >= 256 B
    // when the slabs cut the x axis)
    int best = S; int64_t bestVox = -1; bool bestSlab = false;
    for (int segLen : {32, 64, 128, 256, S}) {
        if (segLen > S) continue;
        std::vector<Box> tmp; boxesFor(segLen, tmp);
        int64_t vox = 0; for (const Box& b : tmp) vox += boxVoxels(b);
        if (bestVox < 0 || vox < bestVox) { bestVox = vox; best = segLen; }
    }
    // slab thickness 64: measured on B200 (C3, end to end): 40 -> 3.22 ms, 48 -> 3.06, 56 -> 3.04, 64 -> 2.82, 96 -> 2.87,
    // 128 -> 2.88, 192 -> 2.94 (thinner slabs move fewer bytes but their short rows starve the copy engine)
    for (int width : {64, 32}) {
        std::vector<Box> tmp;
        if (!slabBoxes(width, tmp)) break;
        int64_t vox = 0; bool narrowRows = false;
        for (const Box& b : tmp) { vox += boxVoxels(b); narrowRows |= (b.hi[0] - b.lo[0] + 1) < 64 && (b.hi[0] - b.lo[0] + 1) < dims[0]; }
        if (width == 32 && narrowRows) continue;     // 128-byte rows starve the copy engine
        if (vox < bestVox) { bestVox = vox; best = width; bestSlab = true; }
    }
    if (bestSlab) slabBoxes(best, out); else boxesFor(best, out);
}

// Dose boxes that enclose every voxel bev_to_grid_kernel can modify for this beam: voxels whose fan index lies within
// one texel of the haloed BEV volume [-32, Rx+32) x [-32, Ry+32) x [bfi, Zg) (trilinear, border 0), clipped to the
// kernel's own coverage. fromFan is affine in each fan index separately -> corners suffice.
void doseBoxesOfBeam(const BeamWs& w, const int dims[3], std::vector<Box>& out) {
    const GridParams& g = w.hostDev.grid;
    if (g.zChunks <= 0) return;
    Box cover;
    cover.lo[0] = g.minX; cover.hi[0] = std::min(g.minX + g.tilesX * 32 - 1, g.nx - 1);
    cover.lo[1] = g.minY; cover.hi[1] = std::min(g.minY + g.tilesY * 8 - 1, g.ny - 1);
    cover.lo[2] = g.minZ; cover.hi[2] = g.maxZ;
    const int Rx = w.grid.dimX, Ry = w.grid.dimY;
    const int segLen = 48;
    for (int k0 = w.bfi; k0 < w.Zg; k0 += segLen) {
        const int k1 = std::min(k0 + segLen, w.Zg - 1);
        float mn[3] = {1e30f, 1e30f, 1e30f}, mx[3] = {-1e30f, -1e30f, -1e30f};
        for (int c = 0; c < 8; ++c) {
            const V3 f = v3((c & 1) ? float(Rx + RTD_MAX_SUPERP_R) : -float(RTD_MAX_SUPERP_R + 1), (c & 2) ? float(Ry + RTD_MAX_SUPERP_R) : -float(RTD_MAX_SUPERP_R + 1),
                            (c & 4) ? float(k1 + 1) : float(k0 - 1));
            const V3 q = fromFan(w.grid.rayIdxToGantry, w.dist, w.gtdi, f);
            const float v[3] = {q.x, q.y, q.z};
            for (int d = 0; d < 3; ++d) { mn[d] = std::min(mn[d], v[d]); mx[d] = std::max(mx[d], v[d]); }
        }
        Box b; bool finite = true;
        for (int d = 0; d < 3; ++d) {
            if (!(std::fabs(mn[d]) < 1e9f && std::fabs(mx[d]) < 1e9f)) { finite = false; break; }
            b.lo[d] = std::max(int(std::floor(mn[d])) - 1, cover.lo[d]); b.hi[d] = std::min(int(std::ceil(mx[d])) + 1, cover.hi[d]);
        }
        if (!finite) b = cover;
        addBox(out, b, dims);
    }
}

// Sub-volume copies: the host / linear-device side is addressed by offsetting the base pointer (positions stay 0), so only
// the array side uses cudaPos, in elements.
cudaError_t copyCtBox(rtd_ctx* ctx, const float* host, const Box& b, cudaStream_t st) {
    const size_t nx = ctx->ctDims[0], ny = ctx->ctDims[1];
    const size_t w = b.hi[0] - b.lo[0] + 1, h = b.hi[1] - b.lo[1] + 1, d = b.hi[2] - b.lo[2] + 1;
    const float* src = host + (size_t(b.lo[2]) * ny + b.lo[1]) * nx + b.lo[0];
    cudaMemcpy3DParms cp; memset(&cp, 0, sizeof cp);
    cp.srcPtr = make_cudaPitchedPtr((void*)src, nx * 4, w, ny);
    cp.dstArray = ctx->ctTex.arr; cp.dstPos = make_cudaPos(b.lo[0], b.lo[1], b.lo[2]);
    cp.extent = make_cudaExtent(w, h, d);          // an array takes part: extent in elements
    cp.kind = cudaMemcpyHostToDevice;
    return cudaMemcpy3DAsync(&cp, st);
}

cudaError_t copyDoseBox(rtd_ctx* ctx, float* host, float* dev, const Box& b, bool toDevice, cudaStream_t st) {
    const size_t nx = ctx->doseDims[0], ny = ctx->doseDims[1];
    const size_t w = b.hi[0] - b.lo[0] + 1, h = b.hi[1] - b.lo[1] + 1, d = b.hi[2] - b.lo[2] + 1;
    const size_t off = (size_t(b.lo[2]) * ny + b.lo[1]) * nx + b.lo[0];
    if (w == nx && h == ny) {   // whole slices: one linear copy
        return toDevice ? cudaMemcpyAsync(dev + off, host + off, d * ny * nx * 4, cudaMemcpyHostToDevice, st)
                        : cudaMemcpyAsync(host + off, dev + off, d * ny * nx * 4, cudaMemcpyDeviceToHost, st);
    }
    cudaMemcpy3DParms cp; memset(&cp, 0, sizeof cp);
    const cudaPitchedPtr hp = make_cudaPitchedPtr((void*)(host + off), nx * 4, w * 4, ny);
    const cudaPitchedPtr dp = make_cudaPitchedPtr((void*)(dev + off), nx * 4, w * 4, ny);
    cp.srcPtr = toDevice ? hp : dp; cp.dstPtr = toDevice ? dp : hp;
    cp.extent = make_cudaExtent(w * 4, h, d);      // no array: width in bytes
    cp.kind = toDevice ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
    return cudaMemcpy3DAsync(&cp, st);
}

// Every beam of a plan is checked before anything is launched: the C ABI (and the --case reader of the host driver) can be handed
// arbitrary bytes, and a non-positive spot pitch or a non-finite sigma would make the `while (dist < lim)` loops of
// spot_to_ray_kernel endless (the reference has the same loops, gpu_convolution_2d.cu:24-33, but only ever sees its own main()).
const char* validateBeam(const rtd_beam& b, const rtd_opts& o) {
    if (b.nLayers <= 0 || b.sx <= 0 || b.sy <= 0 || !b.weights || !b.energies || !b.sigmas) return "Empty list";   // findMax on an empty vector, vector_find.h:22-25
    if (b.steps == 0 || b.steps > 4095) return "beam.steps must be in [1, 4095]";
    if (b.nLayers > 255) return "at most 255 energy layers per beam";
    if (int64_t(b.sx) * b.sy > (1 << 24)) return "spot grid too large";
    auto fin = [](float v) { return std::isfinite(v); };
    for (int i = 0; i < 2; ++i) {
        if (!fin(b.spotDelta[i]) || !(b.spotDelta[i] > 0.0f)) return "beam.spotDelta x / y must be finite and > 0";
        if (!fin(b.raySpacing[i]) || !(b.raySpacing[i] > 0.0f)) return "beam.raySpacing must be finite and > 0";
        if (std::isnan(b.sourceDist[i]) || b.sourceDist[i] == 0.0f) return "beam.sourceDist must be non-zero (INFINITY = parallel beam)";
    }
    if (!fin(b.spotDelta[2]) || b.spotDelta[2] == 0.0f) return "beam.spotDelta z (signed step length) must be finite and non-zero";
    for (int i = 0; i < 3; ++i) if (!fin(b.spotOffset[i]) || !fin(b.gantryToImIdxV[i]) || !fin(b.gantryToDoseIdxV[i])) return "beam offsets / translations must be finite";
    for (int i = 0; i < 9; ++i) if (!fin(b.gantryToImIdxM[i]) || !fin(b.gantryToDoseIdxM[i])) return "beam matrices must be finite";
    for (int l = 0; l < b.nLayers; ++l) {
        if (!fin(b.energies[l])) return "beam.energies must be finite";
        if (!fin(b.sigmas[2 * l]) || !fin(b.sigmas[2 * l + 1]) || !(b.sigmas[2 * l] > 0.0f) || !(b.sigmas[2 * l + 1] > 0.0f)) return "beam.sigmas must be finite and > 0";
    }
    if (!(b.layerBegin == 0 && b.layerEnd == 0) && (b.layerBegin < 0 || b.layerEnd > b.nLayers || b.layerBegin >= b.layerEnd)) return "beam.layerBegin/layerEnd out of range";
    if (!fin(o.convSigmaCutoff) || !(o.convSigmaCutoff > 0.0f) || !fin(o.ksSigmaCutoff) || !(o.ksSigmaCutoff > 0.0f) || !fin(o.bpDepthCutoff) || !fin(o.rayWeightCutoff))
        return "opts cut-offs must be finite (sigma cut-offs > 0)";
    // the pre-convolution visits (2 cutoff sigma / spotDelta) spots per ray and axis: bound it
    float mx = 0.0f, my = 0.0f;
    for (int l = 0; l < b.nLayers; ++l) { mx = std::max(mx, b.sigmas[2 * l]); my = std::max(my, b.sigmas[2 * l + 1]); }
    if (o.convSigmaCutoff * mx / b.spotDelta[0] > 1.0e4f || o.convSigmaCutoff * my / b.spotDelta[1] > 1.0e4f) return "spot sigma is more than 1e4 spot pitches";
    const double nxr = (double(b.sx) * b.spotDelta[0] + 2.0 * o.convSigmaCutoff * mx) / b.raySpacing[0], nyr = (double(b.sy) * b.spotDelta[1] + 2.0 * o.convSigmaCutoff * my) / b.raySpacing[1];
    if (!(nxr < 1.0e5) || !(nyr < 1.0e5) || (nxr / 32.0 + 1.0) * (nyr / 8.0 + 1.0) > 4095.0) return "ray grid has more than 4095 tiles";
    return nullptr;
}

constexpr size_t EV_BEAM0 = 12, EV_PER_BEAM = 14;   // events 0..7 are per call; per beam: 0 begin, 1 traced, 2 set up, 3 conv, 4 idd, 5 radius, 6 superp, 7 transfer, 8 done, 9 CT uploaded, 10 BeamDev on the host
cudaEvent_t evBeam(rtd_ctx* ctx, int b, int i) { return ev(ctx, EV_BEAM0 + size_t(b) * EV_PER_BEAM + i); }

// Part 1 of a beam, everything up to the superposition: launched without knowing any device result (cu:612-1100). The grids
// depend on the geometry alone; cut-offs, slice counts and the transfer box live in BeamDev (beam_setup_kernel).
// Host half: geometry, workspace growth, per-layer host parameters and the staging copies. Nothing is queued here, so a captured
// launch sequence (CUDA graph) can be replayed after it: the graph's copy nodes read the staging buffers this refreshes.
// What the beams of a call that have no (or too small) workspaces yet will ask beamPrepare for, as ONE chunk of the arena.
void reserveWorkspaces(rtd_ctx* ctx, const rtd_beam* beams, int nBeams, const rtd_opts& o) {
    size_t need = 0;
    for (int i = 0; i < nBeams; ++i) {
        const rtd_beam& b = beams[i];
        float mx = b.sigmas[0], my = b.sigmas[1];
        for (int l = 1; l < b.nLayers; ++l) { mx = std::max(mx, b.sigmas[2 * l]); my = std::max(my, b.sigmas[2 * l + 1]); }
        const RayGrid g = makeRayGrid(b.spotDelta, b.spotOffset, b.sx, b.sy, b.raySpacing, mx, my, o.convSigmaCutoff);
        if (g.dimX <= 0 || g.dimY <= 0) continue;
        const size_t Rx = g.dimX, Ry = g.dimY, R = Rx * Ry, S = b.steps, tilesX = Rx / 32, tiles = tilesX * (Ry / 8);
        const size_t L = (b.layerBegin == 0 && b.layerEnd == 0) ? size_t(b.nLayers) : size_t(b.layerEnd - b.layerBegin);
        if (i < (int)ctx->ws.size() && ctx->ws[i]->idd.cap >= L * S * R * 4 && ctx->ws[i]->bev.cap >= S * (Rx + 64) * (Ry + 64) * 4) continue;
        const size_t sum = 2 * R * S * 4 + 2 * L * S * R * 4 + L * S * Ry * tilesX * 4 + 2 * L * R * 4 + L * S * tiles + S * tiles * 4 +
                           S * (Rx + 64) * (Ry + 64) * 4 + L * tilesX * Ry * 2 + size_t(b.sx) * b.sy * L * 4 + 2 * R * 4;
        need += sum + sum / 8 + (size_t(1) << 20);       // ensure() asks for an eighth more; small buffers and alignment in the last term
    }
    if (need) ctx->devArena.reserve(need);
}

rtd_status beamPrepare(rtd_ctx* ctx, const rtd_beam& b, BeamWs& w, const rtd_opts& o) {
    if (brickWanted(o)) { const rtd_status sl = ensureCtLinear(ctx); if (sl != RTD_OK) return sl; }
    float mx = b.sigmas[0], my = b.sigmas[1];
    for (int l = 1; l < b.nLayers; ++l) { mx = std::max(mx, b.sigmas[2 * l]); my = std::max(my, b.sigmas[2 * l + 1]); }
    w.grid = makeRayGrid(b.spotDelta, b.spotOffset, b.sx, b.sy, b.raySpacing, mx, my, o.convSigmaCutoff);
    if (w.grid.dimX <= 0 || w.grid.dimY <= 0) return fail(ctx, RTD_ERR_INVALID, "empty ray grid");
    w.S = (int)b.steps; w.L = b.nLayers; w.R = w.grid.dimX * w.grid.dimY; w.sx = b.sx; w.sy = b.sy;
    if (b.layerBegin == 0 && b.layerEnd == 0) { w.l0 = 0; w.Lc = b.nLayers; }
    else { w.l0 = b.layerBegin; w.Lc = b.layerEnd - b.layerBegin; }
    if ((w.grid.dimX / 32) * (w.grid.dimY / 8) > 4095) return fail(ctx, RTD_ERR_INVALID, "ray grid has more than 4095 tiles");
    w.dist[0] = b.sourceDist[0]; w.dist[1] = b.sourceDist[1];
    w.gtii.m = m3(b.gantryToImIdxM); w.gtii.v = v3(b.gantryToImIdxV[0], b.gantryToImIdxV[1], b.gantryToImIdxV[2]);
    w.gtdi.m = m3(b.gantryToDoseIdxM); w.gtdi.v = v3(b.gantryToDoseIdxV[0], b.gantryToDoseIdxV[1], b.gantryToDoseIdxV[2]);
    const int Rx = w.grid.dimX, Ry = w.grid.dimY, R = w.R, S = w.S, L = w.Lc;
    const int tilesX = Rx / 32, tilesY = Ry / 8, tiles = tilesX * tilesY;

    // ---- workspace, sized for the worst case (every step of the tracer live): nothing here depends on a device result
    const size_t RS = size_t(R) * S, LSR = size_t(L) * S * R;
    RTD_CUDA(w.density.ensure(RS * 4)); RTD_CUDA(w.wepl.ensure(RS * 4));
    RTD_CUDA(w.firstIn.ensure(size_t(R) * 4)); RTD_CUDA(w.firstOut.ensure(size_t(R) * 4));
    RTD_CUDA(w.weplMin.ensure(size_t(S) * 4 + 8));
    const size_t nW = size_t(b.sx) * b.sy * L;
    RTD_CUDA(w.spotW.ensure(nW * 4)); RTD_CUDA(w.hStage.ensure(nW * 4 + size_t(L) * sizeof(LayerIn) + 16));
    RTD_CUDA(w.sigmas.ensure(size_t(L) * 8)); RTD_CUDA(w.layerIn.ensure(size_t(L) * sizeof(LayerIn))); RTD_CUDA(w.layerDev.ensure(size_t(L) * sizeof(LayerParams)));
    RTD_CUDA(w.beamDev.ensure(sizeof(BeamDev))); RTD_CUDA(w.stepTab.ensure(size_t(S) * 8));
    RTD_CUDA(w.rayW.ensure(size_t(L) * R * 4));
    RTD_CUDA(w.idd.ensure(LSR * 4)); RTD_CUDA(w.rsig.ensure(LSR * 4));
    RTD_CUDA(w.rowMin.ensure(size_t(L) * S * Ry * tilesX * 4));
    RTD_CUDA(w.firstPassive.ensure(size_t(L) * R * 4));
    const size_t ctlInts = w.ctlInts();
    RTD_CUDA(w.lfp.ensure(ctlInts * 4));
    RTD_CUDA(w.batchRad.ensure(size_t(L) * 34 * 4)); RTD_CUDA(w.rbOffset.ensure(33 * 4));
    RTD_CUDA(w.radMap.ensure(size_t(L) * S * tiles)); RTD_CUDA(w.list.ensure(size_t(S) * tiles * 4));
    const int planeFloats = (Ry + 64) * (Rx + 64);
    RTD_CUDA(w.bev.ensure(size_t(S) * planeFloats * 4));
    const int segs = tilesX * Ry;                      // 32-ray row segments per layer (< 2^16: at most 4095 tiles of 8 rows)
    RTD_CUDA(w.segList.ensure(size_t(L) * segs * 2));
    RTD_CUDA(w.hOut.ensure(w.outCtlOff() + (ctlInts + size_t(L) * 34) * 4));

    // ---- host part of the per-layer parameters (cu:829-838): needs the LUTs only
    const float* E = ctx->energies.data(); const int nE = ctx->nEnergies;
    float maxEnergy = b.energies[0];
    for (int l = 1; l < w.L; ++l) maxEnergy = std::max(maxEnergy, b.energies[l]);
    const float maxPeakDepth = vectorInterpolate(ctx->peakDepths.data(), nE, findDecimalOrdered(E, nE, maxEnergy));
    w.lin.resize(L);
    for (int l = 0; l < L; ++l) {
        const int al = w.l0 + l;   // absolute layer number in the beam
        LayerIn& li = w.lin[l];
        li.energyIdx = findDecimalOrdered(E, nE, b.energies[al]);
        li.energyScale = vectorInterpolate(ctx->scaleFacts.data(), nE, li.energyIdx);
        li.peakDepth = vectorInterpolate(ctx->peakDepths.data(), nE, li.energyIdx);
        li.thr = o.bpDepthCutoff * li.peakDepth;
        sigmaSqAirCoefsHost(li.peakDepth, !o.noNozzle, li.ca, li.cb);
        li.sigX = b.sigmas[2 * al]; li.sigY = b.sigmas[2 * al + 1];
        li.sigScale = o.nuclearCorr == 3 ? 0.97f : 1.0f;
    }
    memcpy(w.hStage.p, b.weights + size_t(b.sx) * b.sy * w.l0, nW * 4);
    char* stLin = w.hStage.as<char>() + (nW * 4 + 15) / 16 * 16;
    memcpy(stLin, w.lin.data(), size_t(L) * sizeof(LayerIn));
    w.maxPeakDepth = maxPeakDepth;
    return RTD_OK;
}

rtd_status beamEnqueue(rtd_ctx* ctx, int beamNo, const rtd_beam& b, BeamWs& w, const rtd_opts& o, int& launches, cudaStream_t st, HostIo* io) {
    const int Rx = w.grid.dimX, Ry = w.grid.dimY, R = w.R, S = w.S, L = w.Lc;
    const int tilesX = Rx / 32, tilesY = Ry / 8, tiles = tilesX * tilesY;
    const size_t nW = size_t(b.sx) * b.sy * L, ctlInts = w.ctlInts();
    const int planeFloats = (Ry + 64) * (Rx + 64), segs = tilesX * Ry;
    char* stLin = w.hStage.as<char>() + (nW * 4 + 15) / 16 * 16;
    const float maxPeakDepth = w.maxPeakDepth;

    RTD_CUDA(cudaEventRecord(evBeam(ctx, beamNo, 0), st));
    // One-call host path: the beam's small uploads ride on the copy stream in front of its CT boxes. On the beam's own stream they
    // would queue on the copy engine behind EVERY CT box already submitted - those of the beams after this one included - and the
    // beam would not start before the whole plan's CT had arrived (measured: beam 1 of C3 saw its CT 0.31 ms late).
    cudaStream_t up = (io && io->ct) ? ctx->cstream : st;
    RTD_CUDA(cudaMemcpyAsync(w.spotW.p, w.hStage.p, nW * 4, cudaMemcpyHostToDevice, up));
    RTD_CUDA(cudaMemcpyAsync(w.layerIn.p, stLin, size_t(L) * sizeof(LayerIn), cudaMemcpyHostToDevice, up));
    RTD_CUDA(cudaMemsetAsync(w.lfp.p, 0, ctlInts * 4, st));
    const TraceParams tp = makeTraceParams(ctx->densScale, ctx->spScale, b.steps, w.grid.rayIdxToGantry, w.dist, w.gtii);
    if (io && io->ct) {
        // one-call host path: upload the CT sub-volumes this beam's tracer samples (copy stream), then let the beam's stream go
        cudaEvent_t evCt = evBeam(ctx, beamNo, 9);
        if (io->ctMapped) {
            // page-locked CT: the GPU pulls, per CT row, the x interval this beam's tracer can touch and that is not resident yet
            const int rows = ctx->ctDims[1] * ctx->ctDims[2];
            int* need = ctx->ctRows.as<int>();
            RTD_CUDA(cudaMemsetAsync(need, 0x7f, size_t(rows) * 4, ctx->cstream));            // needLo = "nothing"
            RTD_CUDA(cudaMemsetAsync(need + rows, 0xff, size_t(rows) * 4, ctx->cstream));     // needHi = -1
            const long long patches = 2LL * (Rx + Ry) * std::max(S - 1, 1) + 2LL * std::max(Rx - 1, 1) * std::max(Ry - 1, 1);
            ct_mark_rows_kernel<<<(unsigned)((patches + 255) / 256), 256, 0, ctx->cstream>>>(tp, Rx, Ry, S, ctx->ctDims[0], ctx->ctDims[1], ctx->ctDims[2],
                                                                                             need, need + rows);
            ct_pull_rows_kernel<<<(rows + 7) / 8, 256, 0, ctx->cstream>>>(io->ctMapped, ctx->ctTex.surf, ctx->ctDims[0], ctx->ctDims[1], ctx->ctDims[2],
                                                                          need, need + rows, need + 2 * rows, need + 3 * rows,
                                                                          ctx->xferCount.as<unsigned long long>() + 1);
            launches += 2;
            RTD_CUDA(cudaGetLastError());
        } else {
            const size_t n0 = ctx->ctBoxes.size();
            ctBoxesOfBeam(tp, Rx, Ry, S, ctx->ctDims, ctx->ctBoxes);
            for (size_t i = n0; i < ctx->ctBoxes.size(); ++i) {
                RTD_CUDA(copyCtBox(ctx, io->ct, ctx->ctBoxes[i], ctx->cstream));
                io->h2d += boxVoxels(ctx->ctBoxes[i]) * 4;
            }
        }
        RTD_CUDA(cudaEventRecord(evCt, ctx->cstream));
        RTD_CUDA(cudaStreamWaitEvent(st, evCt, 0));
        RTD_CUDA(cudaEventRecord(evBeam(ctx, beamNo, 12), st));      // (timeline: when the beam's stream saw its CT)
    }
    if (io && !io->ct && io->ctReady) RTD_CUDA(cudaStreamWaitEvent(st, evBeam(ctx, beamNo, 9), 0));      // group CT relay (groupCtRelay): this beam's boxes are in
    if (io) io->h2d += int64_t(nW) * 4 + int64_t(L) * sizeof(LayerIn);

    // K1 + reductions. weplMin buffer layout: [0] = beamFirstInside, [1] = beamFirstOutside (ints), [2..] weplMin floats
    int* sc = w.weplMin.as<int>();
    int* dTrFlags = w.lfp.as<int>() + (ctlInts - 4);      // [0] a brick held a non-integer texel, [1] samples outside their brick
    bool brick = brickWanted(o) && ctx->ctLinValid && ctx->ctLin.p && !(io && io->ct);
    BrickParams bp; CUtensorMap tmap;
    if (brick) {
        bp.nx = ctx->ctDims[0]; bp.ny = ctx->ctDims[1]; bp.nz = ctx->ctDims[2]; bp.pitch = ctx->ctPitch; bp.ctLin = ctx->ctLin.as<float>();
        brick = brickParamsOfBeam(tp, Rx, Ry, S, bp) && makeCtTensorMap(ctx, bp, &tmap);
    }
    w.trBrick = brick;
    if (brick) {
        const size_t sh = size_t(bp.BX) * bp.BY * bp.BZ * 4 + 128;
        if (!ctx->brickAttrSet) { RTD_CUDA(cudaFuncSetAttribute(trace_brick_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TRB_SMEM_BUDGET)); ctx->brickAttrSet = true; }
        trace_brick_kernel<<<R / TRB_THREADS, TRB_THREADS, sh, st>>>(tp, tmap, bp, ctx->densTex.tex, ctx->spTex.tex, ctx->densSpTex.tex,
            w.density.as<float>(), w.wepl.as<float>(), w.firstIn.as<int>(), w.firstOut.as<int>(), Rx, R, dTrFlags);
        ++launches;
    }
    // variant A: alone, or - behind variant B - only if that one met a CT texel its hand-made filter cannot reproduce bit for bit
    trace_bev_kernel<<<R / 32, TR_THREADS, 0, st>>>(tp, ctx->ctTex.tex, ctx->densTex.tex, ctx->spTex.tex, ctx->densSpTex.tex,
        w.density.as<float>(), w.wepl.as<float>(), w.firstIn.as<int>(), w.firstOut.as<int>(), Rx, R, brick ? dTrFlags : nullptr);
    slice_reduce_kernel<<<S, 256, 0, st>>>(w.wepl.as<float>(), w.firstIn.as<int>(), w.firstOut.as<int>(), R, w.weplMin.as<float>() + 2, sc, sc + 1);
    RTD_CUDA(cudaEventRecord(evBeam(ctx, beamNo, 1), st));

    // per-beam set-up on the device (the host round trip of cu:781-849)
    SetupIn in;
    in.S = S; in.L = L;
    in.rayResZ = w.grid.rayIdxToGantry.delta.z; in.rayOffZ = w.grid.rayIdxToGantry.offset.z;
    in.distX = w.dist[0]; in.distY = w.dist[1]; in.rayDx = w.grid.rayIdxToGantry.delta.x; in.rayDy = w.grid.rayIdxToGantry.delta.y;
    in.thrAll = o.bpDepthCutoff * maxPeakDepth;
    in.fitf = w.grid.rayIdxToGantry; in.gtdi = w.gtdi; in.dist[0] = w.dist[0]; in.dist[1] = w.dist[1];
    in.t0 = makeTransferParams(w.grid.rayIdxToGantry, w.dist, w.gtdi, v3(float(RTD_MAX_SUPERP_R), float(RTD_MAX_SUPERP_R), 0.0f));
    in.ftfiOffZ = inverse(w.grid.rayIdxToGantry).offset.z;
    in.Rx = Rx; in.Ry = Ry; in.nx = ctx->doseDims[0]; in.ny = ctx->doseDims[1]; in.nz = ctx->doseDims[2];
    in.smCount = ctx->smCount; in.atomic = o.atomicDose ? 1 : 0;
    in.g = makeFillGeom(w.grid.rayIdxToGantry, w.dist);
    BeamDev* bd = w.beamDev.as<BeamDev>();
    beam_setup_kernel<<<1, 256, 0, st>>>(in, sc, w.layerIn.as<LayerIn>(), bd, w.layerDev.as<LayerParams>(), w.sigmas.as<float2>(), w.stepTab.as<float2>());
    // what the host wants to know about the set-up (reports; the dose boxes of a pageable host grid)
    RTD_CUDA(cudaMemcpyAsync(w.hOut.p, w.beamDev.p, sizeof(BeamDev), cudaMemcpyDeviceToHost, st));
    RTD_CUDA(cudaEventRecord(evBeam(ctx, beamNo, 10), st));
    beam_clear_kernel<<<ctx->smCount * 4, 256, 0, st>>>(w.bev.as<float>(), bd, planeFloats, w.rowMin.as<float>(), Ry * tilesX, L, w.firstPassive.as<int>(), L * R);
    RTD_CUDA(cudaEventRecord(evBeam(ctx, beamNo, 2), st));

    int* dLfp = w.lfp.as<int>(); int* dHist = dLfp + L; int* dCursor = dHist + size_t(L) * 34; int* dStats = dCursor + 33; int* dWork = dStats + 4;
    int* dSegCount = dWork + 8; int* dIddQueue = dSegCount + L;
    // K3a (cu:849-854)
    ConvParams cp;
    cp.inW = w.sx; cp.inH = w.sy; cp.outW = Rx; cp.outH = Ry;
    const V3 rayD = w.grid.rayIdxToGantry.delta, rayO = w.grid.rayIdxToGantry.offset;
    cp.deltaX = b.spotDelta[0] / rayD.x; cp.deltaY = b.spotDelta[1] / rayD.y;
    cp.offX = (b.spotOffset[0] - rayO.x) / rayD.x; cp.offY = (b.spotOffset[1] - rayO.y) / rayD.y;
    cp.cutoff = o.convSigmaCutoff; cp.rayWeightCutoff = o.rayWeightCutoff; cp.listAll = o.fullIntermediates ? 1 : 0;
    spot_to_ray_kernel<<<dim3(Rx / 32, L), 256, size_t(w.sy) * 32 * 4, st>>>(w.spotW.as<float>(), w.sigmas.as<float2>(), w.rayW.as<float>(), cp, bd,
                                                                            dSegCount, w.segList.as<uint16_t>());
    RTD_CUDA(cudaEventRecord(evBeam(ctx, beamNo, 3), st));

    // K2
    IddParams q;
    q.g = in.g;
    q.rrlScale = ctx->rrlScale; q.bpCutoff = o.bpDepthCutoff; q.rayWeightCutoff = o.rayWeightCutoff;
    q.Rx = Rx; q.Ry = Ry; q.nozzle = o.noNozzle ? 0 : 1; q.doseToMedium = o.doseToMedium ? 1 : 0; q.writeAll = o.fullIntermediates ? 1 : 0;
    // E_s^2 and the empirical widening per nuclear model (kernel_wrapper.cu:232-245)
    q.nuclear = o.nuclearCorr;
    q.eRefSq = o.nuclearCorr == 1 ? 190.44f : o.nuclearCorr == 2 ? 216.09f : o.nuclearCorr == 3 ? 169.00f : 198.81f;
    q.sigmaDelta = o.nuclearCorr == 1 ? 0.0f : o.nuclearCorr == 2 ? 0.08f : o.nuclearCorr == 3 ? 0.06f : 0.21f;
    // persistent: as many 2-warp CTAs as the machine holds (53 registers -> 18 CTAs per SM), never more than there are segments
    const int iddBlocks = std::max(1, std::min(ctx->iddBlocks, (L * segs + IDD_WARPS - 1) / IDD_WARPS));
    idd_sigma_kernel<<<iddBlocks, IDD_WARPS * 32, 0, st>>>(w.density.as<float>(), w.wepl.as<float>(), w.rayW.as<float>(),
        w.firstIn.as<int>(), w.firstOut.as<int>(), w.layerDev.as<LayerParams>(), bd, w.stepTab.as<float2>(), q, ctx->iddTex.tex, ctx->rrlTex.tex, ctx->nucTex.tex,
        w.idd.as<float>(), w.rsig.as<float>(), w.rowMin.as<float>(), w.firstPassive.as<int>(), dLfp, L, dSegCount, w.segList.as<uint16_t>(), dIddQueue);
    RTD_CUDA(cudaEventRecord(evBeam(ctx, beamNo, 4), st));

    // tile radii, batching, work list (grids for the worst case nz = S; the kernels stop at the beam's own slice count)
    const int gx = (S * tiles + 255) / 256;
    tile_radius_kernel<<<dim3(gx, L), 256, 0, st>>>(w.rowMin.as<float>(), dLfp, bd, Rx, Ry, o.ksSigmaCutoff, q.writeAll ? 0 : 1, w.radMap.as<uint8_t>(), dHist,
                                                     w.idd.as<float>(), w.rsig.as<float>());
    batch_kernel<<<1, 64, 0, st>>>(dHist, L, o.minTilesInBatch, w.batchRad.as<int>(), w.rbOffset.as<int>(), dStats);
    big_items_kernel<<<gx, 256, 0, st>>>(w.radMap.as<uint8_t>(), w.batchRad.as<int>(), L, bd, tiles, w.rbOffset.as<int>() + 16, dCursor, w.list.as<uint32_t>());
    RTD_CUDA(cudaEventRecord(evBeam(ctx, beamNo, 5), st));

    // K3b: layer-fused, warp-autonomous superposition (ks2.cuh), one launch per patch-size class; the heaviest class first
    Ks2Params k2;
    k2.idd = w.idd.as<float>(); k2.rsig = w.rsig.as<float>(); k2.bev = w.bev.as<float>(); k2.batchRad = w.batchRad.as<int>();
    k2.radMap = w.radMap.as<uint8_t>(); k2.layerFirstPassive = dLfp; k2.bd = bd;
    k2.L = L; k2.Rx = Rx; k2.Ry = Ry; k2.tilesX = tilesX; k2.tiles = tiles;
    k2.bigCount = dCursor; k2.bigItems = w.list.as<uint32_t>();
    k2.bodyStep = (o.ks2BodyStep == 1 || o.ks2BodyStep == 2 || o.ks2BodyStep == 4) ? o.ks2BodyStep : 0;
    int* dKsItems = dWork + 4;      // items each radius class processed (rtd_timing.ksLaunchesWithWork)
    // (Cutting the two large classes into several launches on a low-priority stream, so that the tracer .. depth scan of a beam whose CT
    // arrives later on the one-call host path can slip in between, was measured and LOST: the later beam's tracer ends 0.13 ms sooner
    // but every part ends in a tail of idle SMs - C3 end to end 2.55 -> 2.65 (2 parts) -> 2.79 ms (4 parts). One launch per class.)
    ks2_kernel<32, KS2_WARPS_32><<<ctx->ks2Blocks32, KS2_WARPS_32 * 32, KS2_SMEM_32, st>>>(k2, dWork + 0, dKsItems + 0);
    ks2_kernel<24, KS2_WARPS_24><<<ctx->ks2Blocks24, KS2_WARPS_24 * 32, KS2_SMEM_24, st>>>(k2, dWork + 3, dKsItems + 1);
    ks2_kernel<16, KS2_WARPS_16><<<ctx->ks2Blocks16, KS2_WARPS_16 * 32, KS2_SMEM_16, st>>>(k2, dWork + 1, dKsItems + 2);
    ks2_kernel<8, KS2_WARPS_8><<<ctx->ks2Blocks8, KS2_WARPS_8 * 32, KS2_SMEM_8, st>>>(k2, dWork + 2, dKsItems + 3);
    RTD_CUDA(cudaEventRecord(evBeam(ctx, beamNo, 6), st));
    launches += 14;
    RTD_CUDA(cudaGetLastError());
    return RTD_OK;
}

// Part 2 of a beam: the transfer into the dose grid, which all beams share - serialised on the context's main stream in beam
// order, each one waiting for its own beam's superposition (and, on the host path with a pageable grid, for the dose-in
// sub-volumes) - and the copies of the integer results for the reports.
rtd_status beamTransfer(rtd_ctx* ctx, int beamNo, BeamWs& w, float* dDose, int& launches, cudaStream_t st, HostIo* io, cudaEvent_t evDoseIn, bool& doseInWaited) {
    const int L = w.Lc;
    RTD_CUDA(cudaStreamWaitEvent(ctx->stream, evBeam(ctx, beamNo, 6), 0));
    if (io && io->dose && !io->zeroCopy && !doseInWaited) { RTD_CUDA(cudaStreamWaitEvent(ctx->stream, evDoseIn, 0)); doseInWaited = true; }
    bev_to_grid_kernel<<<ctx->smCount * 8, 256, 0, ctx->stream>>>(w.bev.as<float>(), dDose, w.beamDev.as<BeamDev>(), ctx->xferCount.as<unsigned long long>());
    launches += 1;
    RTD_CUDA(cudaGetLastError());
    RTD_CUDA(cudaEventRecord(evBeam(ctx, beamNo, 7), ctx->stream));
    // results for the reports: layer parameters, entry sigmas, control block, batch radii
    char* out = w.hOut.as<char>();
    RTD_CUDA(cudaMemcpyAsync(out + sizeof(BeamDev), w.layerDev.p, size_t(L) * sizeof(LayerParams), cudaMemcpyDeviceToHost, st));
    RTD_CUDA(cudaMemcpyAsync(out + sizeof(BeamDev) + size_t(L) * sizeof(LayerParams), w.sigmas.p, size_t(L) * 8, cudaMemcpyDeviceToHost, st));
    RTD_CUDA(cudaMemcpyAsync(out + w.outCtlOff(), w.lfp.p, w.ctlInts() * 4, cudaMemcpyDeviceToHost, st));
    RTD_CUDA(cudaMemcpyAsync(out + w.outCtlOff() + w.ctlInts() * 4, w.batchRad.p, size_t(L) * 34 * 4, cudaMemcpyDeviceToHost, st));
    if (io) io->d2h += int64_t(sizeof(BeamDev)) + int64_t(L) * (sizeof(LayerParams) + 8) + int64_t(w.ctlInts() + size_t(L) * 34) * 4;
    return RTD_OK;
}

// After the final sync: integer reports (cu:955-957, 1185-1210) and the metric counters.
bool beamFinish(rtd_ctx* ctx, BeamWs& w, int64_t& updates, int64_t& taps, int& ksLaunches) {
    const int L = w.Lc;
    const char* out = w.hOut.as<char>();
    memcpy(&w.hostDev, out, sizeof(BeamDev));
    const LayerParams* lp = reinterpret_cast<const LayerParams*>(out + sizeof(BeamDev));
    const float* esig = reinterpret_cast<const float*>(out + sizeof(BeamDev) + size_t(L) * sizeof(LayerParams));
    w.bfi = w.hostDev.bfi; w.bfo = w.hostDev.bfo; w.Zg = w.hostDev.Zg; w.nz = w.hostDev.nz; w.entryZ = w.hostDev.entryZ;
    w.live = w.nz > 0;
    w.lrep.assign(w.L, rtd_layer_report());
    memset(&w.rep, 0, sizeof w.rep);
    w.rep.lSteps = w.grid.lSteps; w.rep.bSteps = w.grid.bSteps; w.rep.rSteps = w.grid.rSteps; w.rep.tSteps = w.grid.tSteps;
    w.rep.rayDimX = w.grid.dimX; w.rep.rayDimY = w.grid.dimY; w.rep.steps = w.S; w.rep.nLayers = w.L;
    w.rep.beamFirstInside = w.bfi; w.rep.beamFirstOutside = w.bfo; w.rep.beamFirstGuaranteedPassive = w.Zg; w.rep.entryZ = w.entryZ;
    for (int l = 0; l < L; ++l) {
        rtd_layer_report& lr = w.lrep[w.l0 + l];
        lr.energyIdx = w.lin[l].energyIdx; lr.energyScaleFact = w.lin[l].energyScale; lr.peakDepth = w.lin[l].peakDepth;
        lr.entrySigmaX = esig[2 * l]; lr.entrySigmaY = esig[2 * l + 1]; lr.afterLastStep = lp[l].afterLast;
    }
    {   // tracer variant B: did it run, did its result stand, samples it had to filter from global memory
        const int* fl = reinterpret_cast<const int*>(out + w.outCtlOff()) + (w.ctlInts() - 4);
        w.trStood = w.trBrick && fl[0] == 0; w.trOutside = w.trBrick ? fl[1] : 0;
    }
    if (!w.live) return true;
    const int* lfp = reinterpret_cast<const int*>(out + w.outCtlOff()); const int* hist = lfp + L; const int* stats = hist + size_t(L) * 34 + 33;
    const int* ksItems = stats + 8;       // stats[4] | work[4] | ksItems[4] | segCount[L] | iddQueue[4]
    for (int i = 0; i < 4; ++i) if (ksItems[i] > 0) ++ksLaunches;
    const int* brad = lfp + w.ctlInts();
    int fcp = 0; int64_t u = 0, t = 0;
    for (int l = 0; l < L; ++l) {
        rtd_layer_report& lr = w.lrep[w.l0 + l];
        lr.layerFirstPassive = lfp[l];
        fcp = std::max(fcp, lfp[l]);
        const int nLive = std::min(std::max(lfp[l] - w.bfi, 0), w.nz);
        u += int64_t(w.R) * nLive;
        for (int i = 0; i < 34; ++i) {
            lr.radiusHist[i] = hist[l * 34 + i]; lr.batchRadius[i] = brad[l * 34 + i];
            if (i <= 32) { const int64_t rb = brad[l * 34 + i]; t += int64_t(hist[l * 34 + i]) * 256 * (2 * rb + 1) * (2 * rb + 1); }
        }
    }
    w.rep.beamFirstCalculatedPassive = fcp; w.rep.updates = u;
    updates += u; taps += t;
    const int Rx = w.grid.dimX, Ry = w.grid.dimY;
    const float xv[2] = {-float(RTD_MAX_SUPERP_R), float(Rx + RTD_MAX_SUPERP_R - 1)};
    const float yv[2] = {-float(RTD_MAX_SUPERP_R), float(Ry + RTD_MAX_SUPERP_R - 1)};
    const float zv[2] = {float(w.bfi), float(fcp - 1)};
    V3 mx = v3(-1.f, -1.f, -1.f), mn = v3(100000.f, 100000.f, 100000.f);
    for (int k = 0; k < 2; ++k) for (int j = 0; j < 2; ++j) for (int i = 0; i < 2; ++i) {
        const V3 p = fromFan(w.grid.rayIdxToGantry, w.dist, w.gtdi, v3(xv[i], yv[j], zv[k]));
        mx.x = std::max(mx.x, p.x); mx.y = std::max(mx.y, p.y); mx.z = std::max(mx.z, p.z);
        mn.x = std::min(mn.x, p.x); mn.y = std::min(mn.y, p.y); mn.z = std::min(mn.z, p.z);
    }
    w.rep.bboxMin[0] = std::max((int(std::floor(mn.x)) / 32) * 32, 0); w.rep.bboxMin[1] = std::max(int(std::floor(mn.y)), 0); w.rep.bboxMin[2] = std::max(int(std::floor(mn.z)), 0);
    w.rep.bboxMax[0] = std::min(int(std::ceil(mx.x)), ctx->doseDims[0] - 1); w.rep.bboxMax[1] = std::min(int(std::ceil(mx.y)), ctx->doseDims[1] - 1);
    w.rep.bboxMax[2] = std::min(int(std::ceil(mx.z)), ctx->doseDims[2] - 1);
    return stats[0] == 0;
}

// Everything of a call up to (not including) the final host synchronisation.
rtd_status doseEnqueue(rtd_ctx* ctx, const rtd_beam* beams, int nBeams, float* dDose, const rtd_opts& o,
                       cudaEvent_t evBegin, cudaEvent_t evEnd, HostIo* io, CallState& cs) {
    int& launches = cs.launches;
    for (int b = 0; b < nBeams; ++b)
        if (const char* why = validateBeam(beams[b], o)) return fail(ctx, RTD_ERR_INVALID, why);
    if (o.nuclearCorr < 0 || o.nuclearCorr > 3) return fail(ctx, RTD_ERR_INVALID, "opts.nuclearCorr must be 0 (off), 1 (SOUKUP), 2 (FLUKA) or 3 (GAUSS_FIT)");
    if (o.nuclearCorr && !ctx->nucTex.tex) return fail(ctx, RTD_ERR_STATE, "opts.nuclearCorr needs rtd_luts.nucWeightMatrix (rtd_set_luts)");
    while ((int)ctx->ws.size() < nBeams) { ctx->ws.push_back(new BeamWs()); ctx->ws.back()->adopt(&ctx->devArena, &ctx->pinArena); };
    ctx->nBeamsLast = nBeams; ctx->optsLast = o;
    ctx->ctBoxes.clear(); ctx->doseBoxes.clear();
    RTD_CUDA(ctx->xferCount.ensure(16)); RTD_CUDA(ctx->hXferCount.ensure(16));   // [0] dose voxels modified, [1] CT bytes pulled
    if (io && io->ctMapped) RTD_CUDA(ctx->ctRows.ensure(size_t(ctx->ctDims[1]) * ctx->ctDims[2] * 16));
    RTD_CUDA(cudaMemsetAsync(ctx->xferCount.p, 0, 16, ctx->stream));
    RTD_CUDA(cudaEventRecord(evBegin, ctx->stream));
    if (io) RTD_CUDA(cudaStreamWaitEvent(ctx->cstream, evBegin, 0));
    if (io && io->ctMapped) {   // nothing of the caller's CT is resident at the start of a call (its content may have changed)
        const size_t rows = size_t(ctx->ctDims[1]) * ctx->ctDims[2];
        RTD_CUDA(cudaMemsetAsync(ctx->ctRows.as<int>() + 2 * rows, 0x7f, rows * 4, ctx->cstream));
        RTD_CUDA(cudaMemsetAsync(ctx->ctRows.as<int>() + 3 * rows, 0xff, rows * 4, ctx->cstream));
    }
    // Beams are independent until the transfer into the shared dose grid: unless opts.serialBeams is set every beam runs
    // on its own stream so that the latency-bound stages (tracer, IDD/sigma depth scans) of different beams overlap; on the
    // one-call host path the CT sub-volumes of beam b+1 upload (copy stream) while beam b computes.
    const bool serial = o.serialBeams != 0 || nBeams == 1;
    cs.serial = serial;
    if (!cs.prepared) {
        reserveWorkspaces(ctx, beams, nBeams, o);
        for (int b = 0; b < nBeams; ++b) { rtd_status s = beamPrepare(ctx, beams[b], *ctx->ws[b], o); if (s != RTD_OK) return s; }
    }
    auto beamStream = [&](int b) -> cudaStream_t {
        if (serial) return ctx->stream;
        const size_t i = size_t(b) % 8;
        while (ctx->bstreams.size() <= i) { cudaStream_t s_; cudaStreamCreateWithFlags(&s_, cudaStreamNonBlocking); ctx->bstreams.push_back(s_); }
        return ctx->bstreams[i];
    };
    int& hostSyncs = cs.hostSyncs;
    cudaEvent_t evDoseIn = ev(ctx, 6);
    const bool stagedDose = io && io->dose && !io->zeroCopy;      // pageable host grid: the transfers need its sub-volumes on the device first
    bool doseInWaited = false;
    for (int b = 0; b < nBeams; ++b) {
        cudaStream_t st = beamStream(b);
        if (st != ctx->stream) RTD_CUDA(cudaStreamWaitEvent(st, evBegin, 0));
        rtd_status s = beamEnqueue(ctx, b, beams[b], *ctx->ws[b], o, launches, st, io);
        if (s != RTD_OK) return s;
        if (!stagedDose) {
            s = beamTransfer(ctx, b, *ctx->ws[b], dDose, launches, st, io, evDoseIn, doseInWaited);
            if (s != RTD_OK) return s;
            if (st != ctx->stream) { cudaEvent_t e = evBeam(ctx, b, 8); RTD_CUDA(cudaEventRecord(e, st)); RTD_CUDA(cudaStreamWaitEvent(ctx->stream, e, 0)); }
        }
    }
    if (stagedDose) {
        // pageable host dose grid: the sub-volumes the transfers can touch are staged through device memory, and their boxes
        // need the transfer boxes on the host - the one case that still waits for the set-up kernels in the middle of a call.
        // Every box is uploaded before the first transfer, so overlapping boxes of different beams are harmless.
        for (int b = 0; b < nBeams; ++b) { RTD_CUDA(cudaEventSynchronize(evBeam(ctx, b, 10))); memcpy(&ctx->ws[b]->hostDev, ctx->ws[b]->hOut.p, sizeof(BeamDev)); }
        hostSyncs = 2;
        for (int b = 0; b < nBeams; ++b) { BeamWs& w = *ctx->ws[b]; w.bfi = w.hostDev.bfi; w.Zg = w.hostDev.Zg; doseBoxesOfBeam(w, ctx->doseDims, ctx->doseBoxes); }
        for (const Box& bx : ctx->doseBoxes) { RTD_CUDA(copyDoseBox(ctx, io->dose, dDose, bx, true, ctx->cstream)); io->h2d += boxVoxels(bx) * 4; }
        RTD_CUDA(cudaEventRecord(evDoseIn, ctx->cstream));
        for (int b = 0; b < nBeams; ++b) {
            cudaStream_t st = beamStream(b);
            rtd_status s = beamTransfer(ctx, b, *ctx->ws[b], dDose, launches, st, io, evDoseIn, doseInWaited);
            if (s != RTD_OK) return s;
            if (st != ctx->stream) { cudaEvent_t e = evBeam(ctx, b, 8); RTD_CUDA(cudaEventRecord(e, st)); RTD_CUDA(cudaStreamWaitEvent(ctx->stream, e, 0)); }
        }
    }
    cudaEvent_t evXferEnd = ev(ctx, 7);
    cs.evXferEnd = evXferEnd;
    RTD_CUDA(cudaEventRecord(evXferEnd, ctx->stream));
    RTD_CUDA(cudaMemcpyAsync(ctx->hXferCount.p, ctx->xferCount.p, 16, cudaMemcpyDeviceToHost, ctx->stream));
    if (io && io->dose && !io->zeroCopy) {
        if (!doseInWaited) RTD_CUDA(cudaStreamWaitEvent(ctx->stream, evDoseIn, 0));
        for (const Box& bx : ctx->doseBoxes) { RTD_CUDA(copyDoseBox(ctx, io->dose, dDose, bx, false, ctx->stream)); io->d2h += boxVoxels(bx) * 4; }
    }
    RTD_CUDA(cudaEventRecord(evEnd, ctx->stream));
    return RTD_OK;
}

// After the call's one host synchronisation: reports, counters, timing.
rtd_status doseFinish(rtd_ctx* ctx, int nBeams, rtd_timing* tm, cudaEvent_t evBegin, cudaEvent_t evEnd, HostIo* io, const CallState& cs) {
    const int launches = cs.launches, hostSyncs = cs.hostSyncs;
    cudaEvent_t evXferEnd = cs.evXferEnd;
    int64_t updates = 0, taps = 0; bool ok = true; int ksLaunches = 0;
    for (int b = 0; b < nBeams; ++b) ok = beamFinish(ctx, *ctx->ws[b], updates, taps, ksLaunches) && ok;
    const int64_t modified = (int64_t)*ctx->hXferCount.as<unsigned long long>();
    if (io && io->ctMapped) io->h2d += (int64_t)ctx->hXferCount.as<unsigned long long>()[1];   // CT bytes the pull kernels read over PCIe
    if (io && io->dose && io->zeroCopy) { io->h2d += modified * 4; io->d2h += modified * 4; }   // the transfers read-modify-write the caller's pinned grid in place
    if (!cs.graph && getenv("RTD_TIMELINE")) {      // debugging aid: when did every stage of every beam END, in ms after the call's first event
        for (int b = 0; b < nBeams; ++b) {
            static const char* name[13] = {"begin", "traced", "setup", "conv", "idd", "radius", "superp", "transfer", "-", "ct", "-", "-", "ctseen"};
            fprintf(stderr, "[rtd timeline] beam %d:", b);
            for (int i : {0, 9, 12, 1, 2, 3, 4, 5, 6, 7}) {
                float ms = -1.0f;
                if ((i == 9 && !(io && (io->ct || io->ctReady))) || (i == 12 && !(io && io->ct))) continue;
                if (cudaEventElapsedTime(&ms, evBegin, evBeam(ctx, b, i)) != cudaSuccess) { cudaGetLastError(); continue; }
                fprintf(stderr, " %s %.3f", name[i], ms);
            }
            fprintf(stderr, "\n");
        }
    }
    if (tm) {
        memset(tm, 0, sizeof *tm);
        tm->updates = updates; tm->ksTaps = taps; tm->kernelLaunches = launches; tm->hostSyncs = hostSyncs; tm->ksLaunchesWithWork = ksLaunches;
        for (int b = 0; b < nBeams; ++b) { tm->tracerBrickBeams += ctx->ws[b]->trStood ? 1 : 0; tm->tracerOutsideSamples += ctx->ws[b]->trOutside; }
        float ms = 0;
        for (int b = 0; b < nBeams && !cs.graph; ++b) {      // (a replayed graph does not record the stage events: stage times stay 0)
            cudaEvent_t e[8];
            for (int i = 0; i < 8; ++i) e[i] = evBeam(ctx, b, i);
            cudaEventElapsedTime(&ms, e[0], e[1]); tm->msTrace += ms;
            cudaEventElapsedTime(&ms, e[1], e[2]); tm->msSetup += ms;
            cudaEventElapsedTime(&ms, e[2], e[3]); tm->msConv += ms;
            cudaEventElapsedTime(&ms, e[3], e[4]); tm->msIddSigma += ms;
            cudaEventElapsedTime(&ms, e[4], e[5]); tm->msRadius += ms;
            cudaEventElapsedTime(&ms, e[5], e[6]); tm->msSuperp += ms;
            cudaEventElapsedTime(&ms, e[6], e[7]); tm->msTransfer += std::max(ms, 0.0f);
        }
        cudaEventElapsedTime(&ms, evBegin, evEnd); tm->msTotal = ms;
        if (io) {
            if (io->ct && nBeams > 0) { cudaEventElapsedTime(&ms, evBegin, evBeam(ctx, nBeams - 1, 9)); tm->msUpload = ms; }
            if (io->dose && !cs.graph) { cudaEventElapsedTime(&ms, evXferEnd, evEnd); tm->msDownload = ms; }
            tm->h2dBytes = io->h2d; tm->d2hBytes = io->d2h;
        }
        tm->doseVoxelsModified = modified;
    }
    if (!ok) return fail(ctx, RTD_ERR_RADIUS, "Found larger than allowed kernel superposition radius");
    if (ctx->optsLast.nuclearCorr)
        for (int b = 0; b < nBeams; ++b)
            if (ctx->ws[b]->live && ctx->ws[b]->bfi == 0)
                return fail(ctx, RTD_ERR_INVALID, "opts.nuclearCorr: a beam starts inside the patient (beamFirstInside == 0); the reference then deposits the "
                                                  "slice-0 residue of its halo pass (kernel_wrapper.cu:925), which is not reproduced - the dose of this call lacks it");
    return RTD_OK;
}

// 64-bit FNV-1a over everything a captured launch sequence bakes in: the beams' scalars and pointers, the arrays the host parameters
// are derived from (energies, sigmas), the options, the dose pointer, the LUT / CT generation and the workspace generation.
uint64_t planKey(rtd_ctx* ctx, const rtd_beam* beams, int nBeams, const float* dDose, const rtd_opts& o) {
    uint64_t h = 1469598103934665603ull;
    auto mix = [&](const void* p, size_t n) { const unsigned char* c = (const unsigned char*)p; for (size_t i = 0; i < n; ++i) { h ^= c[i]; h *= 1099511628211ull; } };
    mix(&nBeams, sizeof nBeams); mix(&dDose, sizeof dDose); mix(&o, sizeof o);
    mix(ctx->ctDims, sizeof ctx->ctDims); mix(ctx->doseDims, sizeof ctx->doseDims); mix(&ctx->ctTex.tex, sizeof ctx->ctTex.tex);
    mix(&ctx->lutGeneration, sizeof ctx->lutGeneration); mix(&ctx->ctLinValid, sizeof ctx->ctLinValid);
    const long re = DevBuf::reallocs(); mix(&re, sizeof re);
    for (int b = 0; b < nBeams; ++b) {
        mix(&beams[b], sizeof(rtd_beam));
        mix(beams[b].energies, size_t(beams[b].nLayers) * 4); mix(beams[b].sigmas, size_t(beams[b].nLayers) * 8);
    }
    return h;
}

// Device-resident plans with several beams are captured once into a CUDA graph (the ~15 launches, copies and cross-stream waits per
// beam become one launch) and replayed for as long as nothing the capture baked in changes; anything else runs eagerly.
rtd_status doseGraph(rtd_ctx* ctx, const rtd_beam* beams, int nBeams, float* dDose, const rtd_opts& o, rtd_timing* tm, bool& handled) {
    handled = false;
    while ((int)ctx->ws.size() < nBeams) { ctx->ws.push_back(new BeamWs()); ctx->ws.back()->adopt(&ctx->devArena, &ctx->pinArena); };
    for (int b = 0; b < nBeams; ++b)
        if (const char* why = validateBeam(beams[b], o)) return fail(ctx, RTD_ERR_INVALID, why);
    reserveWorkspaces(ctx, beams, nBeams, o);
    for (int b = 0; b < nBeams; ++b) { rtd_status s = beamPrepare(ctx, beams[b], *ctx->ws[b], o); if (s != RTD_OK) return s; }
    RTD_CUDA(ctx->xferCount.ensure(16)); RTD_CUDA(ctx->hXferCount.ensure(16));
    while (ctx->bstreams.size() < size_t(std::min(nBeams, 8))) { cudaStream_t s_; RTD_CUDA(cudaStreamCreateWithFlags(&s_, cudaStreamNonBlocking)); ctx->bstreams.push_back(s_); }
    for (int b = 0; b < nBeams; ++b) evBeam(ctx, b, int(EV_PER_BEAM) - 1);      // events exist before the capture starts
    const uint64_t key = planKey(ctx, beams, nBeams, dDose, o);
    cudaEvent_t evB = ev(ctx, 4), evE = ev(ctx, 5);
    if (!ctx->graphExec || ctx->graphKey != key) {
        if (ctx->graphExec) { cudaGraphExecDestroy(ctx->graphExec); ctx->graphExec = nullptr; }
        CallState cs; cs.prepared = true;
        if (cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed) != cudaSuccess) { cudaGetLastError(); return RTD_OK; }
        const rtd_status s = doseEnqueue(ctx, beams, nBeams, dDose, o, ev(ctx, 0), ev(ctx, 1), nullptr, cs);
        cudaGraph_t g = nullptr;
        const cudaError_t e = cudaStreamEndCapture(ctx->stream, &g);
        if (s != RTD_OK || e != cudaSuccess || !g) { if (g) cudaGraphDestroy(g); cudaGetLastError(); ctx->graphBroken = true; return s; }
        const cudaError_t ei = cudaGraphInstantiate(&ctx->graphExec, g, 0);
        cudaGraphDestroy(g);
        if (ei != cudaSuccess) { ctx->graphExec = nullptr; cudaGetLastError(); ctx->graphBroken = true; return RTD_OK; }
        ctx->graphKey = key; ctx->graphCs = cs; ctx->graphCs.graph = true;
        ctx->nBeamsLast = nBeams; ctx->optsLast = o;
    }
    handled = true;
    RTD_CUDA(cudaEventRecord(evB, ctx->stream));
    RTD_CUDA(cudaGraphLaunch(ctx->graphExec, ctx->stream));
    RTD_CUDA(cudaEventRecord(evE, ctx->stream));
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->nBeamsLast = nBeams; ctx->optsLast = o;
    return doseFinish(ctx, nBeams, tm, evB, evE, nullptr, ctx->graphCs);
}

rtd_status doseCoreImpl(rtd_ctx* ctx, const rtd_beam* beams, int nBeams, float* dDose, const rtd_opts& o, rtd_timing* tm,
                        cudaEvent_t evBegin, cudaEvent_t evEnd, HostIo* io) {
    if (!io && nBeams > 1 && !o.serialBeams && !o.noGraph && !ctx->graphBroken) {
        bool handled = false;
        const rtd_status gs = doseGraph(ctx, beams, nBeams, dDose, o, tm, handled);
        if (gs != RTD_OK || handled) return gs;
    }
    CallState cs;
    rtd_status s = doseEnqueue(ctx, beams, nBeams, dDose, o, evBegin, evEnd, io, cs);
    if (s != RTD_OK) return s;
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));                         // the one host sync of a plan
    return doseFinish(ctx, nBeams, tm, evBegin, evEnd, io, cs);
}

// A failure in the middle of a plan leaves work queued on the per-beam streams and the copy stream: drain them all before the
// status reaches the caller, who may free or reuse its buffers right away.
rtd_status doseCore(rtd_ctx* ctx, const rtd_beam* beams, int nBeams, float* dDose, const rtd_opts& o, rtd_timing* tm,
                    cudaEvent_t evBegin, cudaEvent_t evEnd, HostIo* io) {
    const rtd_status s = doseCoreImpl(ctx, beams, nBeams, dDose, o, tm, evBegin, evEnd, io);
    if (s != RTD_OK) {
        const std::string msg = ctx->err;
        for (cudaStream_t bs : ctx->bstreams) cudaStreamSynchronize(bs);
        cudaStreamSynchronize(ctx->cstream); cudaStreamSynchronize(ctx->stream);
        cudaGetLastError();
        ctx->err = msg;
    }
    return s;
}

}  // namespace

// =========================================================================================================
extern "C" {

const char* rtd_version(void) { return "rtdose-b200 0.1 (sm_100a)"; }

void rtd_default_opts(rtd_opts* o) {
    if (!o) return;
    memset(o, 0, sizeof *o);
    o->bpDepthCutoff = 1.05f; o->convSigmaCutoff = 3.0f; o->ksSigmaCutoff = 3.0f; o->rayWeightCutoff = 1.0f;
    o->minTilesInBatch = 16; o->tracerBrick = -1;
}

static rtd_status createCommon(rtd_ctx** out, int device, cudaStream_t userStream, bool haveUserStream) {
    if (!out) return RTD_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) return RTD_ERR_NO_DEVICE;   // no CPU fallback by design
    if (device < 0 || device >= n) return RTD_ERR_INVALID;
    if (cudaSetDevice(device) != cudaSuccess) return RTD_ERR_CUDA;
    rtd_ctx* ctx = new rtd_ctx();
    ctx->device = device;
    if (haveUserStream) { ctx->stream = userStream; ctx->ownStream = false; }
    else if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { delete ctx; return RTD_ERR_CUDA; }
    int prLo = 0, prHi = 0; cudaDeviceGetStreamPriorityRange(&prLo, &prHi);   // the copy stream also runs the CT pull kernels: highest priority
    if (cudaStreamCreateWithPriority(&ctx->cstream, cudaStreamNonBlocking, prHi) != cudaSuccess) { if (ctx->ownStream) cudaStreamDestroy(ctx->stream); delete ctx; return RTD_ERR_CUDA; }
    cudaDeviceGetAttribute(&ctx->smCount, cudaDevAttrMultiProcessorCount, device);
    rtd_default_opts(&ctx->optsLast);
    for (DevBuf* b : {&ctx->ctRows, &ctx->xferCount, &ctx->dose, &ctx->ctLin, &ctx->groupEpochDev}) b->arena = &ctx->devArena;
    for (PinBuf* b : {&ctx->hXferCount, &ctx->groupFlag}) b->arena = &ctx->pinArena;
    cudaFuncSetAttribute(ks2_kernel<16, KS2_WARPS_16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)KS2_SMEM_16);
    cudaFuncSetAttribute(ks2_kernel<8, KS2_WARPS_8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)KS2_SMEM_8);
    cudaFuncSetAttribute(ks2_kernel<32, KS2_WARPS_32>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)KS2_SMEM_32);
    cudaFuncSetAttribute(ks2_kernel<24, KS2_WARPS_24>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)KS2_SMEM_24);
    int occ = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ks2_kernel<16, KS2_WARPS_16>, KS2_WARPS_16 * 32, KS2_SMEM_16); ctx->ks2Blocks16 = std::max(occ, 1) * ctx->smCount;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ks2_kernel<8, KS2_WARPS_8>, KS2_WARPS_8 * 32, KS2_SMEM_8); ctx->ks2Blocks8 = std::max(occ, 1) * ctx->smCount;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ks2_kernel<32, KS2_WARPS_32>, KS2_WARPS_32 * 32, KS2_SMEM_32); ctx->ks2Blocks32 = std::max(occ, 1) * ctx->smCount;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, ks2_kernel<24, KS2_WARPS_24>, KS2_WARPS_24 * 32, KS2_SMEM_24); ctx->ks2Blocks24 = std::max(occ, 1) * ctx->smCount;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, idd_sigma_kernel, IDD_WARPS * 32, 0); ctx->iddBlocks = std::max(occ, 1) * ctx->smCount;
    if (cudaGetLastError() != cudaSuccess) { cudaStreamDestroy(ctx->cstream); if (ctx->ownStream) cudaStreamDestroy(ctx->stream); delete ctx; return RTD_ERR_CUDA; }
    *out = ctx;
    return RTD_OK;
}

rtd_status rtd_create(rtd_ctx** out, int device) { return createCommon(out, device, nullptr, false); }
rtd_status rtd_create_on_stream(rtd_ctx** out, int device, void* cudaStream) { return createCommon(out, device, (cudaStream_t)cudaStream, true); }

void rtd_destroy(rtd_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    for (BeamWs* w : ctx->ws) { w->release(); delete w; }
    ctx->iddTex.release(); ctx->densTex.release(); ctx->spTex.release(); ctx->densSpTex.release(); ctx->rrlTex.release(); ctx->ctTex.release();
    ctx->nucTex.release();
    ctx->dose.release(); ctx->xferCount.release(); ctx->hXferCount.release(); ctx->ctRows.release(); ctx->ctLin.release();
    if (ctx->graphExec) cudaGraphExecDestroy(ctx->graphExec);
    if (ctx->groupGrid) rtd_group_destroy(ctx);
    ctx->groupEpochDev.release(); ctx->groupFlag.release();
    ctx->devArena.release(); ctx->pinArena.release();
    for (cudaEvent_t e : ctx->events) cudaEventDestroy(e);
    for (cudaStream_t s_ : ctx->bstreams) { cudaStreamSynchronize(s_); cudaStreamDestroy(s_); }
    cudaStreamSynchronize(ctx->cstream); cudaStreamDestroy(ctx->cstream);
    if (ctx->ownStream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char* rtd_last_error(const rtd_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

void* rtd_stream(rtd_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }

rtd_status rtd_set_luts(rtd_ctx* ctx, const rtd_luts* l) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!l || l->nEnergies < 2 || l->nEnergySamples < 2 || !l->energiesPerU || !l->peakDepths || !l->scaleFacts || !l->ciddMatrix ||
        l->nDensitySamples < 1 || !l->densityVector || l->nSpSamples < 1 || !l->spVector || l->nRRlSamples < 1 || !l->rRlVector)
        return fail(ctx, RTD_ERR_INVALID, "rtd_set_luts: incomplete LUT set");
    RTD_CUDA(cudaSetDevice(ctx->device));
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->nEnergies = l->nEnergies; ctx->nEnergySamples = l->nEnergySamples;
    ctx->energies.assign(l->energiesPerU, l->energiesPerU + l->nEnergies);
    ctx->peakDepths.assign(l->peakDepths, l->peakDepths + l->nEnergies);
    ctx->scaleFacts.assign(l->scaleFacts, l->scaleFacts + l->nEnergies);
    ctx->densScale = l->densityScaleFact; ctx->spScale = l->spScaleFact; ctx->rrlScale = l->rRlScaleFact;
    ctx->iddTex.release();
    cudaChannelFormatDesc fd = cudaCreateChannelDesc<float>();
    RTD_CUDA(cudaMallocArray(&ctx->iddTex.arr, &fd, l->nEnergySamples, l->nEnergies));
    RTD_CUDA(cudaMemcpy2DToArrayAsync(ctx->iddTex.arr, 0, 0, l->ciddMatrix, size_t(l->nEnergySamples) * 4, size_t(l->nEnergySamples) * 4,
                                      l->nEnergies, cudaMemcpyHostToDevice, ctx->stream));
    RTD_CUDA(makeTexture(ctx->iddTex, cudaAddressModeClamp));
    ctx->nucTex.release();
    if (l->nucWeightMatrix) {   // kernel_wrapper.cu:544-566: same shape and addressing as the IDD matrix
        RTD_CUDA(cudaMallocArray(&ctx->nucTex.arr, &fd, l->nEnergySamples, l->nEnergies));
        RTD_CUDA(cudaMemcpy2DToArrayAsync(ctx->nucTex.arr, 0, 0, l->nucWeightMatrix, size_t(l->nEnergySamples) * 4, size_t(l->nEnergySamples) * 4,
                                          l->nEnergies, cudaMemcpyHostToDevice, ctx->stream));
        RTD_CUDA(makeTexture(ctx->nucTex, cudaAddressModeClamp));
    }
    RTD_CUDA(upload1D(ctx->rrlTex, l->rRlVector, l->nRRlSamples, ctx->stream));
    ctx->densSpTex.release(); ctx->densTex.release(); ctx->spTex.release();
    std::vector<float2> both;
    const bool oneLut = l->nDensitySamples == l->nSpSamples && l->densityScaleFact == l->spScaleFact;
    if (!oneLut) {      // (with one float2 LUT the tracers never touch the two single ones: two arrays less to create and free - 0.3 ms each way apiece)
        RTD_CUDA(upload1D(ctx->densTex, l->densityVector, l->nDensitySamples, ctx->stream));
        RTD_CUDA(upload1D(ctx->spTex, l->spVector, l->nSpSamples, ctx->stream));
    }
    if (oneLut) {
        both.resize(l->nDensitySamples);
        for (int i = 0; i < l->nDensitySamples; ++i) both[i] = make_float2(l->densityVector[i], l->spVector[i]);
        cudaChannelFormatDesc fd = cudaCreateChannelDesc<float2>();
        RTD_CUDA(cudaMallocArray(&ctx->densSpTex.arr, &fd, both.size()));
        RTD_CUDA(cudaMemcpy2DToArrayAsync(ctx->densSpTex.arr, 0, 0, both.data(), both.size() * 8, both.size() * 8, 1, cudaMemcpyHostToDevice, ctx->stream));
        RTD_CUDA(makeTexture(ctx->densSpTex, cudaAddressModeClamp));
    }
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->haveLuts = true; ++ctx->lutGeneration;
    return RTD_OK;
}

rtd_status rtd_set_ct(rtd_ctx* ctx, const float* hu, int nx, int ny, int nz) { return setCtCommon(ctx, hu, nx, ny, nz, cudaMemcpyHostToDevice); }
rtd_status rtd_set_ct_device(rtd_ctx* ctx, const float* hu, int nx, int ny, int nz) { return setCtCommon(ctx, hu, nx, ny, nz, cudaMemcpyDeviceToDevice); }

rtd_status rtd_set_dose_grid(rtd_ctx* ctx, int nx, int ny, int nz) {
    if (!ctx) return RTD_ERR_INVALID;
    if (nx <= 0 || ny <= 0 || nz <= 0) return fail(ctx, RTD_ERR_INVALID, "rtd_set_dose_grid: non-positive dims");
    ctx->doseDims[0] = nx; ctx->doseDims[1] = ny; ctx->doseDims[2] = nz; ctx->haveDose = true;
    return RTD_OK;
}

static rtd_status checkReady(rtd_ctx* ctx, const rtd_beam* beams, int nBeams, const void* dose) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!ctx->haveLuts || !ctx->haveCt || !ctx->haveDose) return fail(ctx, RTD_ERR_STATE, "rtd_dose: LUTs, CT and dose grid must be set first");
    if (ctx->ctPartial) return fail(ctx, RTD_ERR_STATE, "rtd_dose: the CT is only partially resident after rtd_plan_dose; call rtd_set_ct again");
    if (nBeams < 0 || (nBeams > 0 && !beams) || !dose) return fail(ctx, RTD_ERR_INVALID, "rtd_dose: bad beams / dose pointer");
    return RTD_OK;
}

rtd_status rtd_dose_device(rtd_ctx* ctx, const rtd_beam* beams, int nBeams, float* dDose, const rtd_opts* opts, rtd_timing* tm) {
    rtd_status s = checkReady(ctx, beams, nBeams, dDose);
    if (s != RTD_OK) return s;
    RTD_CUDA(cudaSetDevice(ctx->device));
    rtd_opts o; if (opts) o = *opts; else rtd_default_opts(&o);
    return doseCore(ctx, beams, nBeams, dDose, o, tm, ev(ctx, 0), ev(ctx, 1), nullptr);
}

// Host dose buffer: only the sub-volumes the beams can modify travel (see rtd_plan_dose in rtdose.h).
static rtd_status doseHost(rtd_ctx* ctx, const float* ct, const rtd_beam* beams, int nBeams, float* dose, float* dDose, const rtd_opts* opts, rtd_timing* tm) {
    rtd_opts o; if (opts) o = *opts; else rtd_default_opts(&o);
    HostIo io; io.ct = ct; io.dose = dose;
    if (ct && o.ctPull) {
        const size_t n = size_t(ctx->ctDims[0]) * ctx->ctDims[1] * ctx->ctDims[2];
        cudaPointerAttributes a0, a1;
        if (cudaPointerGetAttributes(&a0, ct) == cudaSuccess && cudaPointerGetAttributes(&a1, ct + (n - 1)) == cudaSuccess &&
            a0.type == cudaMemoryTypeHost && a1.type == cudaMemoryTypeHost && a0.devicePointer &&
            (char*)a1.devicePointer - (char*)a0.devicePointer == (ptrdiff_t)((n - 1) * 4) && (ctx->ctDims[0] % 32) == 0 &&
            ((uintptr_t)a0.devicePointer & 127) == 0)
            io.ctMapped = (const float*)a0.devicePointer;
        cudaGetLastError();
    }
    if (dose) {
        const size_t n = size_t(ctx->doseDims[0]) * ctx->doseDims[1] * ctx->doseDims[2];
        // Pinned (cudaHostAlloc / cudaHostRegister, as HostPinnedImage3D does, host_image_3d.cuh:24) dose grids are
        // accumulated IN PLACE over PCIe by the transfer kernels: only voxels that receive dose are read and written,
        // in the reference's summation order ((dose + beam0) + beam1 ...). Pageable grids go through sub-volume copies.
        float* mapped = nullptr;
        if (!o.noZeroCopyDose) {
            cudaPointerAttributes a0, a1;
            if (cudaPointerGetAttributes(&a0, dose) == cudaSuccess && cudaPointerGetAttributes(&a1, dose + (n - 1)) == cudaSuccess &&
                a0.type == cudaMemoryTypeHost && a1.type == cudaMemoryTypeHost && a0.devicePointer &&
                (char*)a1.devicePointer - (char*)a0.devicePointer == (ptrdiff_t)((n - 1) * 4))
                mapped = (float*)a0.devicePointer;
            cudaGetLastError();
        }
        if (mapped) { io.zeroCopy = true; dDose = mapped; }
        else { RTD_CUDA(ctx->dose.ensure(n * 4)); dDose = ctx->dose.as<float>(); }
    }
    return doseCore(ctx, beams, nBeams, dDose, o, tm, ev(ctx, 0), ev(ctx, 1), &io);
}

rtd_status rtd_dose(rtd_ctx* ctx, const rtd_beam* beams, int nBeams, float* dose, const rtd_opts* opts, rtd_timing* tm) {
    rtd_status s = checkReady(ctx, beams, nBeams, dose);
    if (s != RTD_OK) return s;
    RTD_CUDA(cudaSetDevice(ctx->device));
    return doseHost(ctx, nullptr, beams, nBeams, dose, nullptr, opts, tm);
}

static rtd_status planDose(rtd_ctx* ctx, const float* ct, int nx, int ny, int nz, const rtd_beam* beams, int nBeams, float* dose, float* dDose,
                           const rtd_opts* opts, rtd_timing* tm) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!ct || nx <= 0 || ny <= 0 || nz <= 0) return fail(ctx, RTD_ERR_INVALID, "rtd_plan_dose: null CT pointer or non-positive dims");
    if (!ctx->haveLuts || !ctx->haveDose) return fail(ctx, RTD_ERR_STATE, "rtd_plan_dose: LUTs and dose grid must be set first");
    if (nBeams < 0 || (nBeams > 0 && !beams) || (!dose && !dDose)) return fail(ctx, RTD_ERR_INVALID, "rtd_plan_dose: bad beams / dose pointer");
    RTD_CUDA(cudaSetDevice(ctx->device));
    rtd_status s = ensureCtArray(ctx, nx, ny, nz);
    if (s != RTD_OK) return s;
    ctx->haveCt = true; ctx->ctPartial = true; ctx->ctLinValid = false;
    return doseHost(ctx, ct, beams, nBeams, dose, dDose, opts, tm);
}

rtd_status rtd_plan_dose(rtd_ctx* ctx, const float* ct, int nx, int ny, int nz, const rtd_beam* beams, int nBeams, float* dose,
                         const rtd_opts* opts, rtd_timing* tm) {
    return planDose(ctx, ct, nx, ny, nz, beams, nBeams, dose, nullptr, opts, tm);
}
rtd_status rtd_plan_dose_device(rtd_ctx* ctx, const float* ct, int nx, int ny, int nz, const rtd_beam* beams, int nBeams, float* dDose,
                                const rtd_opts* opts, rtd_timing* tm) {
    return planDose(ctx, ct, nx, ny, nz, beams, nBeams, nullptr, dDose, opts, tm);
}

// ---- one dose grid shared by the ranks of a node (CUDA IPC over NVLink / NVSwitch) -------------------------
rtd_status rtd_shared_dose_create(rtd_ctx* ctx, size_t nFloats, float** dPtr, unsigned char* handle64) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!dPtr || !handle64 || nFloats == 0) return fail(ctx, RTD_ERR_INVALID, "rtd_shared_dose_create: bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    RTD_CUDA(cudaSetDevice(ctx->device));
    void* p = nullptr;
    RTD_CUDA(cudaMalloc(&p, nFloats * 4));
    cudaIpcMemHandle_t h;
    cudaError_t e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); char b[256]; snprintf(b, sizeof b, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e)); return fail(ctx, RTD_ERR_CUDA, b); }
    RTD_CUDA(cudaMemsetAsync(p, 0, nFloats * 4, ctx->stream));
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    memcpy(handle64, &h, 64);
    *dPtr = (float*)p;
    return RTD_OK;
}

rtd_status rtd_shared_dose_open(rtd_ctx* ctx, const unsigned char* handle64, float** dPtr) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!dPtr || !handle64) return fail(ctx, RTD_ERR_INVALID, "rtd_shared_dose_open: bad arguments");
    RTD_CUDA(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h; memcpy(&h, handle64, 64);
    void* p = nullptr;
    RTD_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *dPtr = (float*)p;
    return RTD_OK;
}

rtd_status rtd_shared_dose_close(rtd_ctx* ctx, float* dPtr, int owner) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!dPtr) return RTD_OK;
    RTD_CUDA(cudaSetDevice(ctx->device));
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    if (owner) RTD_CUDA(cudaFree(dPtr)); else RTD_CUDA(cudaIpcCloseMemHandle(dPtr));
    return RTD_OK;
}

rtd_status rtd_add_device_to_host(rtd_ctx* ctx, const float* dSrc, float* hostDst, size_t n, int64_t* modifiedOut) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!dSrc || !hostDst) return fail(ctx, RTD_ERR_INVALID, "rtd_add_device_to_host: null pointer");
    RTD_CUDA(cudaSetDevice(ctx->device));
    if (modifiedOut) *modifiedOut = 0;
    if (n == 0) return RTD_OK;
    cudaPointerAttributes a0, a1;
    float* mapped = nullptr;
    if (cudaPointerGetAttributes(&a0, hostDst) == cudaSuccess && cudaPointerGetAttributes(&a1, hostDst + (n - 1)) == cudaSuccess &&
        a0.type == cudaMemoryTypeHost && a1.type == cudaMemoryTypeHost && a0.devicePointer &&
        (char*)a1.devicePointer - (char*)a0.devicePointer == (ptrdiff_t)((n - 1) * 4))
        mapped = (float*)a0.devicePointer;
    cudaGetLastError();
    if (mapped) {
        RTD_CUDA(ctx->xferCount.ensure(8)); RTD_CUDA(ctx->hXferCount.ensure(8));
        RTD_CUDA(cudaMemsetAsync(ctx->xferCount.p, 0, 8, ctx->stream));
        const int blocks = (int)std::min<size_t>((n + 255) / 256, size_t(ctx->smCount) * 16);
        add_nonzero_kernel<<<blocks, 256, 0, ctx->stream>>>(dSrc, mapped, n, ctx->xferCount.as<unsigned long long>());
        RTD_CUDA(cudaGetLastError());
        RTD_CUDA(cudaMemcpyAsync(ctx->hXferCount.p, ctx->xferCount.p, 8, cudaMemcpyDeviceToHost, ctx->stream));
        RTD_CUDA(cudaStreamSynchronize(ctx->stream));
        if (modifiedOut) *modifiedOut = (int64_t)*ctx->hXferCount.as<unsigned long long>();
    } else {   // pageable destination: stage through host memory
        std::vector<float> tmp(n);
        RTD_CUDA(cudaMemcpyAsync(tmp.data(), dSrc, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
        RTD_CUDA(cudaStreamSynchronize(ctx->stream));
        int64_t m = 0;
        for (size_t i = 0; i < n; ++i) if (tmp[i] != 0.0f) { hostDst[i] += tmp[i]; ++m; }
        if (modifiedOut) *modifiedOut = m;
    }
    return RTD_OK;
}

rtd_status rtd_push_dose_to_peer(rtd_ctx* ctx, const float* dSrc, float* dPeerDst, size_t n) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!dSrc || !dPeerDst) return fail(ctx, RTD_ERR_INVALID, "rtd_push_dose_to_peer: null pointer");
    if (((uintptr_t)dSrc | (uintptr_t)dPeerDst) & 15) return fail(ctx, RTD_ERR_INVALID, "rtd_push_dose_to_peer: pointers must be 16-byte aligned");
    RTD_CUDA(cudaSetDevice(ctx->device));
    const size_t n4 = n / 4;
    if (n4) {
        const int blocks = (int)std::min<size_t>((n4 + 255) / 256, size_t(ctx->smCount) * 16);
        push_nonzero_kernel<<<blocks, 256, 0, ctx->stream>>>((const float4*)dSrc, (float4*)dPeerDst, n4);
    }
    if (n & 3) push_nonzero_tail_kernel<<<1, 32, 0, ctx->stream>>>(dSrc + n4 * 4, dPeerDst + n4 * 4, (int)(n & 3));
    RTD_CUDA(cudaGetLastError());
    return RTD_OK;
}

// ---- one plan on the GPUs of a node: partition, barriers and the sum of the partial doses behind the C ABI ---------------
int rtd_partition(const rtd_beam* beams, int nBeams, const rtd_luts* luts, int world, int chunksPerBeam, rtd_work_item* items, int maxItems) {
    if (nBeams < 0 || (nBeams > 0 && !beams) || !luts || !luts->energiesPerU || !luts->peakDepths || luts->nEnergies < 2 || world < 1) return -1;
    for (int b = 0; b < nBeams; ++b) if (beams[b].nLayers <= 0 || !beams[b].energies) return -1;
    const std::vector<WorkItem> v = partitionPlan(beams, nBeams, luts->energiesPerU, luts->peakDepths, luts->nEnergies, world, chunksPerBeam);
    for (size_t i = 0; i < v.size() && (int)i < maxItems && items; ++i) { items[i].rank = v[i].rank; items[i].beam = v[i].beam; items[i].layerBegin = v[i].lo; items[i].layerEnd = v[i].hi; }
    return (int)v.size();
}

int rtd_dose_slab(const rtd_beam* beams, int nBeams, int nx, int ny, int nz, float convSigmaCutoff, int* z0, int* z1) {
    if (nBeams < 0 || (nBeams > 0 && !beams) || nx <= 0 || ny <= 0 || nz <= 0 || !z0 || !z1) return -1;
    for (int b = 0; b < nBeams; ++b) if (beams[b].nLayers <= 0 || !beams[b].sigmas) return -1;
    const int dims[3] = {nx, ny, nz};
    return doseSlab(beams, nBeams, dims, convSigmaCutoff, *z0, *z1) ? 1 : 0;
}

static void groupCtRelayClose(rtd_ctx* ctx);
rtd_status rtd_group_destroy(rtd_ctx* ctx) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!ctx->groupGrid) return RTD_OK;
    RTD_CUDA(cudaSetDevice(ctx->device));
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    groupCtRelayClose(ctx);
    if (ctx->groupRank == 0) RTD_CUDA(cudaFree(ctx->groupGrid)); else RTD_CUDA(cudaIpcCloseMemHandle(ctx->groupGrid));
    if (ctx->groupGraphExec) { cudaGraphExecDestroy(ctx->groupGraphExec); ctx->groupGraphExec = nullptr; }
    ctx->groupGrid = nullptr; ctx->groupCtr = nullptr; ctx->groupRank = -1; ctx->groupWorld = 0; ctx->groupFloats = 0;
    return RTD_OK;
}

rtd_status rtd_group_create(rtd_ctx* ctx, int rank, int world, size_t nDoseFloats, unsigned char* handle64) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!handle64 || world < 1 || rank < 0 || rank >= world || nDoseFloats == 0) return fail(ctx, RTD_ERR_INVALID, "rtd_group_create: bad arguments");
    rtd_status s = rtd_group_destroy(ctx);
    if (s != RTD_OK) return s;
    RTD_CUDA(cudaSetDevice(ctx->device));
    const size_t gridBytes = (nDoseFloats * 4 + 255) / 256 * 256;
    void* p = nullptr;
    if (rank == 0) {
        RTD_CUDA(cudaMalloc(&p, gridBytes + 256));
        cudaIpcMemHandle_t h;
        cudaError_t e = cudaSuccess;
        if (world > 1) e = cudaIpcGetMemHandle(&h, p); else memset(&h, 0, sizeof h);
        if (e != cudaSuccess) { cudaFree(p); char b[256]; snprintf(b, sizeof b, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e)); return fail(ctx, RTD_ERR_CUDA, b); }
        RTD_CUDA(cudaMemsetAsync(p, 0, gridBytes + 256, ctx->stream));
        RTD_CUDA(cudaStreamSynchronize(ctx->stream));
        memcpy(handle64, &h, 64);
    } else {
        cudaIpcMemHandle_t h; memcpy(&h, handle64, 64);
        RTD_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    }
    RTD_CUDA(ctx->groupFlag.ensure(64));
    memset(ctx->groupFlag.p, 0, 64);
    ctx->groupGrid = (float*)p; ctx->groupCtr = (unsigned long long*)((char*)p + gridBytes);
    RTD_CUDA(ctx->groupEpochDev.ensure(8));
    RTD_CUDA(cudaMemsetAsync(ctx->groupEpochDev.p, 0, 8, ctx->stream));
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->groupRank = rank; ctx->groupWorld = world; ctx->groupFloats = nDoseFloats;
    return RTD_OK;
}

float* rtd_group_dose(rtd_ctx* ctx) { return ctx ? ctx->groupGrid : nullptr; }

static rtd_status groupBarrierEnqueue(rtd_ctx* ctx) {
    if (ctx->groupWorld <= 1) return RTD_OK;
    group_barrier_kernel<<<1, 1, 0, ctx->stream>>>(ctx->groupCtr, ctx->groupEpochDev.as<unsigned long long>(), (unsigned long long)ctx->groupWorld,
                                                   ctx->groupFlag.as<int>(), 10ull * 1000 * 1000 * 1000);
    RTD_CUDA(cudaGetLastError());
    return RTD_OK;
}

rtd_status rtd_group_barrier(rtd_ctx* ctx) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!ctx->groupGrid) return fail(ctx, RTD_ERR_STATE, "rtd_group_barrier: no group (rtd_group_create)");
    RTD_CUDA(cudaSetDevice(ctx->device));
    return groupBarrierEnqueue(ctx);
}

// ---- CT relay of a group ("CT upload at N = 8") -------------------------------------------
// Without it every rank uploads the CT sub-volumes of every beam it holds a layer of over its own PCIe link: at N = 8 the node moves
// 168 MB for a plan whose union of sub-volumes is 61 MB, and a rank with pieces of two beams waits for 40 MB before its second
// tracer can start. With it the plan's boxes are dealt to the ranks by bytes, each rank uploads its share into its own staging buffer,
// one group barrier later every rank copies the boxes its beams touch from the owners' buffers into its CT array - device to device
// over NVLink / NVSwitch peer memory.
static void groupCtRelayClose(rtd_ctx* ctx) {
    for (size_t r = 0; r < ctx->ctRelayPeers.size(); ++r)
        if (ctx->ctRelayPeers[r] && (int)r != ctx->groupRank) cudaIpcCloseMemHandle(ctx->ctRelayPeers[r]);
    ctx->ctRelayPeers.clear();
    if (ctx->ctRelay) cudaFree(ctx->ctRelay);
    ctx->ctRelay = nullptr; ctx->ctRelayBytes = 0; ctx->ctRelayOpen = false;
    cudaGetLastError();
}

rtd_status rtd_group_ct_create(rtd_ctx* ctx, size_t bytes, unsigned char* handle64) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!ctx->groupGrid) return fail(ctx, RTD_ERR_STATE, "rtd_group_ct_create: no group (rtd_group_create)");
    if (!handle64 || bytes < 4096) return fail(ctx, RTD_ERR_INVALID, "rtd_group_ct_create: bad arguments");
    RTD_CUDA(cudaSetDevice(ctx->device));
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    groupCtRelayClose(ctx);
    RTD_CUDA(cudaMalloc(&ctx->ctRelay, bytes));
    ctx->ctRelayBytes = bytes;
    cudaIpcMemHandle_t h; memset(&h, 0, sizeof h);
    if (ctx->groupWorld > 1) {
        const cudaError_t e = cudaIpcGetMemHandle(&h, ctx->ctRelay);
        if (e != cudaSuccess) { groupCtRelayClose(ctx); char b[256]; snprintf(b, sizeof b, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e)); return fail(ctx, RTD_ERR_CUDA, b); }
    }
    memcpy(handle64, &h, 64);
    return RTD_OK;
}

rtd_status rtd_group_ct_open(rtd_ctx* ctx, const unsigned char* handles) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!ctx->groupGrid || !ctx->ctRelay) return fail(ctx, RTD_ERR_STATE, "rtd_group_ct_open: rtd_group_create and rtd_group_ct_create come first");
    if (!handles) return fail(ctx, RTD_ERR_INVALID, "rtd_group_ct_open: null handles");
    RTD_CUDA(cudaSetDevice(ctx->device));
    ctx->ctRelayPeers.assign(size_t(ctx->groupWorld), nullptr);
    for (int r = 0; r < ctx->groupWorld; ++r) {
        if (r == ctx->groupRank) { ctx->ctRelayPeers[r] = ctx->ctRelay; continue; }
        cudaIpcMemHandle_t h; memcpy(&h, handles + size_t(r) * 64, 64);
        const cudaError_t e = cudaIpcOpenMemHandle(&ctx->ctRelayPeers[r], h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            ctx->ctRelayPeers[r] = nullptr; groupCtRelayClose(ctx);
            char b[256]; snprintf(b, sizeof b, "cudaIpcOpenMemHandle (CT relay of rank %d): %s", r, cudaGetErrorString(e)); return fail(ctx, RTD_ERR_CUDA, b);
        }
    }
    ctx->ctRelayOpen = true;
    return RTD_OK;
}

// Ray grid and tracer geometry of a beam as beamPrepare derives them (nothing is allocated)
static bool beamCtGeometry(const rtd_beam& b, const rtd_opts& o, TraceParams& tp, int& Rx, int& Ry) {
    float mx = b.sigmas[0], my = b.sigmas[1];
    for (int l = 1; l < b.nLayers; ++l) { mx = std::max(mx, b.sigmas[2 * l]); my = std::max(my, b.sigmas[2 * l + 1]); }
    const RayGrid grid = makeRayGrid(b.spotDelta, b.spotOffset, b.sx, b.sy, b.raySpacing, mx, my, o.convSigmaCutoff);
    if (grid.dimX <= 0 || grid.dimY <= 0) return false;
    Affine gtii; gtii.m = m3(b.gantryToImIdxM); gtii.v = v3(b.gantryToImIdxV[0], b.gantryToImIdxV[1], b.gantryToImIdxV[2]);
    tp = makeTraceParams(1.0f, 1.0f, b.steps, grid.rayIdxToGantry, b.sourceDist, gtii);
    Rx = grid.dimX; Ry = grid.dimY;
    return true;
}

// The relay phase of one call. `mine[k]` = plan beam of this rank's k-th beam (the order they are enqueued in). On success with `used` = true the CT array holds
// every texel this rank's tracers can fetch once io.ctReady has passed. `used` stays false (and nothing is enqueued) if the relay is not
// open or a rank's share would not fit its staging buffer - a pure function of the plan, so every rank decides the same.
static rtd_status groupCtRelay(rtd_ctx* ctx, const float* ct, const rtd_beam* beams, int nBeams, const std::vector<int>& mine, const rtd_opts& o,
                               HostIo& io, bool& used) {
    used = false;
    const int world = ctx->groupWorld, rank = ctx->groupRank;
    if (!ctx->ctRelayOpen || world <= 1) return RTD_OK;
    const int* dims = ctx->ctDims;
    std::vector<Box> U;                       // the plan's boxes, each minus what earlier ones hold (addBox): a partition of their union
    std::vector<std::vector<Box>> need;       // per beam of this rank, in its order: what the tracer can touch (the beam's boxes, not reduced)
    for (int b = 0; b < nBeams; ++b) {
        TraceParams tp; int Rx = 0, Ry = 0;
        if (!beamCtGeometry(beams[b], o, tp, Rx, Ry)) return fail(ctx, RTD_ERR_INVALID, "empty ray grid");
        ctBoxesOfBeam(tp, Rx, Ry, (int)beams[b].steps, dims, U);
    }
    for (size_t k = 0; k < mine.size(); ++k) {
        TraceParams tp; int Rx = 0, Ry = 0;
        if (!beamCtGeometry(beams[mine[k]], o, tp, Rx, Ry)) return fail(ctx, RTD_ERR_INVALID, "empty ray grid");
        need.emplace_back(); ctBoxesOfBeam(tp, Rx, Ry, (int)beams[mine[k]].steps, dims, need.back());
    }
    int64_t total = 0;
    for (const Box& u : U) total += boxVoxels(u) * 4;
    if (total == 0) return RTD_OK;
    // boxes are dealt out in list order by bytes: owner = the rank whose byte range holds the box's first byte
    std::vector<int> owner(U.size()); std::vector<size_t> offset(U.size());
    std::vector<int64_t> share(size_t(world), 0);
    int64_t cum = 0;
    for (size_t i = 0; i < U.size(); ++i) {
        const int r = (int)std::min<int64_t>(cum * world / total, world - 1);
        owner[i] = r; offset[i] = size_t(share[r]); share[r] += (boxVoxels(U[i]) * 4 + 255) / 256 * 256;
        cum += boxVoxels(U[i]) * 4;
    }
    for (int r = 0; r < world; ++r) if (size_t(share[r]) > ctx->ctRelayBytes) return RTD_OK;      // (every rank was created with the same size)
    const size_t nx = dims[0], ny = dims[1];
    cudaEvent_t evUp = ev(ctx, 8), evBar = ev(ctx, 9), evReady = ev(ctx, 10);
    // 1. this rank's share: host -> own staging buffer, every box packed densely
    for (size_t i = 0; i < U.size(); ++i) {
        if (owner[i] != rank) continue;
        const Box& u = U[i];
        const size_t w = u.hi[0] - u.lo[0] + 1, h = u.hi[1] - u.lo[1] + 1, d = u.hi[2] - u.lo[2] + 1;
        cudaMemcpy3DParms cp; memset(&cp, 0, sizeof cp);
        cp.srcPtr = make_cudaPitchedPtr((void*)(ct + (size_t(u.lo[2]) * ny + u.lo[1]) * nx + u.lo[0]), nx * 4, w * 4, ny);
        cp.dstPtr = make_cudaPitchedPtr((char*)ctx->ctRelay + offset[i], w * 4, w * 4, h);
        cp.extent = make_cudaExtent(w * 4, h, d); cp.kind = cudaMemcpyHostToDevice;
        RTD_CUDA(cudaMemcpy3DAsync(&cp, ctx->cstream));
        io.h2d += boxVoxels(u) * 4;
    }
    // 2. every share has landed (stream-ordered barrier of the group, no host sync)
    RTD_CUDA(cudaEventRecord(evUp, ctx->cstream));
    RTD_CUDA(cudaStreamWaitEvent(ctx->stream, evUp, 0));
    { const rtd_status s = groupBarrierEnqueue(ctx); if (s != RTD_OK) return s; }
    RTD_CUDA(cudaEventRecord(evBar, ctx->stream));
    RTD_CUDA(cudaStreamWaitEvent(ctx->cstream, evBar, 0));
    // 3. the boxes this rank's beams touch: owner's staging buffer (peer memory) -> own CT array, beam by beam in the order the beams are
    //    enqueued, each with its own event (the slot the per-rank upload uses): the first tracer starts after its own boxes
    std::vector<char> pulled(U.size(), 0);
    for (size_t k = 0; k < need.size(); ++k) {
        for (size_t i = 0; i < U.size(); ++i) {
            if (pulled[i]) continue;
            const Box& u = U[i];
            bool hit = false;
            for (const Box& n : need[k]) {
                bool x = true;
                for (int d = 0; d < 3; ++d) x = x && u.lo[d] <= n.hi[d] && n.lo[d] <= u.hi[d];
                if (x) { hit = true; break; }
            }
            if (!hit) continue;
            pulled[i] = 1;
            const size_t w = u.hi[0] - u.lo[0] + 1, h = u.hi[1] - u.lo[1] + 1, d = u.hi[2] - u.lo[2] + 1;
            // (copy engine, 77 GB/s into the array. A kernel writing the boxes through the surface is faster alone but
            // queues behind the superposition launches of the beams already running, which own every SM: beam 2 of C4 on two GPUs saw
            // its CT 4 ms late. Measured, not used.)
            cudaMemcpy3DParms cp; memset(&cp, 0, sizeof cp);
            cp.srcPtr = make_cudaPitchedPtr((char*)ctx->ctRelayPeers[owner[i]] + offset[i], w * 4, w, h);
            cp.dstArray = ctx->ctTex.arr; cp.dstPos = make_cudaPos(u.lo[0], u.lo[1], u.lo[2]);
            cp.extent = make_cudaExtent(w, h, d); cp.kind = cudaMemcpyDeviceToDevice;
            RTD_CUDA(cudaMemcpy3DAsync(&cp, ctx->cstream));
        }
        RTD_CUDA(cudaGetLastError());
        RTD_CUDA(cudaEventRecord(evBeam(ctx, (int)k, 9), ctx->cstream));
    }
    RTD_CUDA(cudaEventRecord(evReady, ctx->cstream));
    io.ctReady = evReady;
    used = true;
    return RTD_OK;
}

rtd_status rtd_plan_dose_group(rtd_ctx* ctx, const float* ct, int nx, int ny, int nz, const rtd_beam* beams, int nBeams, float* hostDoseInOut,
                               const rtd_opts* opts, rtd_timing* tm) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!ctx->groupGrid) return fail(ctx, RTD_ERR_STATE, "rtd_plan_dose_group: no group (rtd_group_create)");
    if (!ctx->haveLuts || !ctx->haveDose) return fail(ctx, RTD_ERR_STATE, "rtd_plan_dose_group: LUTs and dose grid must be set first");
    if (nBeams < 0 || (nBeams > 0 && !beams)) return fail(ctx, RTD_ERR_INVALID, "rtd_plan_dose_group: bad beams");
    const size_t nGrid = size_t(ctx->doseDims[0]) * ctx->doseDims[1] * ctx->doseDims[2];
    if (nGrid != ctx->groupFloats) return fail(ctx, RTD_ERR_INVALID, "rtd_plan_dose_group: the group's grid does not have the dose grid's size");
    RTD_CUDA(cudaSetDevice(ctx->device));
    rtd_opts o; if (opts) o = *opts; else rtd_default_opts(&o);
    for (int b = 0; b < nBeams; ++b)
        if (const char* why = validateBeam(beams[b], o)) return fail(ctx, RTD_ERR_INVALID, why);
    if (ct) {
        if (nx <= 0 || ny <= 0 || nz <= 0) return fail(ctx, RTD_ERR_INVALID, "rtd_plan_dose_group: non-positive CT dims");
        rtd_status s = ensureCtArray(ctx, nx, ny, nz);
        if (s != RTD_OK) return s;
        ctx->haveCt = true; ctx->ctPartial = true; ctx->ctLinValid = false;
    } else {
        if (!ctx->haveCt || ctx->ctPartial) return fail(ctx, RTD_ERR_STATE, "rtd_plan_dose_group: no CT given and no complete CT resident (rtd_set_ct)");
    }
    const int rank = ctx->groupRank, world = ctx->groupWorld;
    // this rank's work items: the same partition on every rank (a pure function of the plan)
    const std::vector<WorkItem> items = partitionPlan(beams, nBeams, ctx->energies.data(), ctx->peakDepths.data(), ctx->nEnergies, world, 0);
    ctx->groupBeams.clear();
    for (const WorkItem& it : items) {
        if (it.rank != rank) continue;
        rtd_beam b = beams[it.beam];
        if (!(it.lo == 0 && it.hi == b.nLayers)) { b.layerBegin = it.lo; b.layerEnd = it.hi; } else { b.layerBegin = 0; b.layerEnd = 0; }
        ctx->groupBeams.push_back(b);
    }
    int z0 = 0, z1 = -1;
    const bool haveSlab = doseSlab(beams, nBeams, ctx->doseDims, o.convSigmaCutoff, z0, z1);
    const size_t sliceFloats = size_t(ctx->doseDims[0]) * ctx->doseDims[1];
    const size_t slabOff = haveSlab ? size_t(z0) * sliceFloats : 0, slabN = haveSlab ? size_t(z1 - z0 + 1) * sliceFloats : 0;
    float* target = ctx->groupGrid;
    if (rank != 0) { RTD_CUDA(ctx->dose.ensure(nGrid * 4)); target = ctx->dose.as<float>(); }
    cudaEvent_t evBegin = ev(ctx, 0), evEnd = ev(ctx, 1), evAll = ev(ctx, 3), evStart = ev(ctx, 4);
    // rank 0 computes straight into the plan's grid while the pushes of the others may land in it: its transfers add atomically
    o.atomicDose = (rank == 0 && world > 1) ? 1 : o.atomicDose;
    HostIo io; io.ct = ct;
    CallState cs;
    const int nMine = (int)ctx->groupBeams.size();
    if (ct) {      // CT relay (if the caller opened one): the plan's CT boxes cross PCIe once on the node
        std::vector<int> mine;
        for (const WorkItem& it : items) if (it.rank == rank) mine.push_back(it.beam);
        bool relayed = false;
        const rtd_status r = groupCtRelay(ctx, ct, beams, nBeams, mine, o, io, relayed);
        if (r != RTD_OK) return r;
        if (relayed) io.ct = nullptr;
    }
    // zero the slab | barrier | this rank's beams | push | barrier: the whole device-side sequence of a rank
    auto enqueueSeq = [&](CallState& c) -> rtd_status {
        if (slabN) RTD_CUDA(cudaMemsetAsync(target + slabOff, 0, slabN * 4, ctx->stream));
        rtd_status r = groupBarrierEnqueue(ctx);             // the plan's grid is zeroed before anything is added into it
        if (r == RTD_OK) r = doseEnqueue(ctx, ctx->groupBeams.data(), nMine, target, o, evBegin, evEnd, ct ? &io : nullptr, c);
        if (r == RTD_OK && rank != 0 && slabN) r = rtd_push_dose_to_peer(ctx, target + slabOff, ctx->groupGrid + slabOff, slabN);
        if (r == RTD_OK) r = groupBarrierEnqueue(ctx);       // every rank's contribution has landed
        return r;
    };
    rtd_status s = RTD_OK;
    bool replayed = false;
    if (!ct && !hostDoseInOut && !o.noGraph && !o.serialBeams && !ctx->graphBroken) {
        // device-resident share of the plan: captured once, replayed while nothing the capture baked in changes (see doseGraph)
        reserveWorkspaces(ctx, ctx->groupBeams.data(), nMine, o);
        for (int b = 0; b < nMine && s == RTD_OK; ++b) { while ((int)ctx->ws.size() < nMine) { ctx->ws.push_back(new BeamWs()); ctx->ws.back()->adopt(&ctx->devArena, &ctx->pinArena); }; s = beamPrepare(ctx, ctx->groupBeams[b], *ctx->ws[b], o); }
        if (s != RTD_OK) return s;
        RTD_CUDA(ctx->xferCount.ensure(16)); RTD_CUDA(ctx->hXferCount.ensure(16));
        while (ctx->bstreams.size() < size_t(std::min(std::max(nMine, 1), 8))) { cudaStream_t s_; RTD_CUDA(cudaStreamCreateWithFlags(&s_, cudaStreamNonBlocking)); ctx->bstreams.push_back(s_); }
        for (int b = 0; b < std::max(nMine, 1); ++b) evBeam(ctx, b, int(EV_PER_BEAM) - 1);
        uint64_t key = planKey(ctx, ctx->groupBeams.data(), nMine, target, o);
        key ^= (uint64_t(rank) << 48) ^ (uint64_t(world) << 56) ^ uint64_t(slabOff) * 1099511628211ull ^ uint64_t(slabN) ^ uint64_t(uintptr_t(ctx->groupGrid));
        if (!ctx->groupGraphExec || ctx->groupGraphKey != key) {
            if (ctx->groupGraphExec) { cudaGraphExecDestroy(ctx->groupGraphExec); ctx->groupGraphExec = nullptr; }
            CallState c; c.prepared = true;
            if (cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeRelaxed) == cudaSuccess) {
                const rtd_status r = enqueueSeq(c);
                cudaGraph_t g = nullptr;
                const cudaError_t e = cudaStreamEndCapture(ctx->stream, &g);
                if (r == RTD_OK && e == cudaSuccess && g && cudaGraphInstantiate(&ctx->groupGraphExec, g, 0) == cudaSuccess) { ctx->groupGraphKey = key; ctx->graphCs = c; ctx->graphCs.graph = true; }
                else { ctx->groupGraphExec = nullptr; ctx->graphBroken = true; cudaGetLastError(); }
                if (g) cudaGraphDestroy(g);
                if (r != RTD_OK) return r;
            } else cudaGetLastError();
        }
        if (ctx->groupGraphExec) {
            RTD_CUDA(cudaEventRecord(evStart, ctx->stream));
            RTD_CUDA(cudaGraphLaunch(ctx->groupGraphExec, ctx->stream));
            cs = ctx->graphCs; replayed = true;
            ctx->nBeamsLast = nMine; ctx->optsLast = o;
        }
    }
    if (!replayed) {
        RTD_CUDA(cudaEventRecord(evStart, ctx->stream));
        s = enqueueSeq(cs);
    }
    int64_t modified = 0;
    if (s == RTD_OK && rank == 0 && hostDoseInOut && slabN) {
        // the summed slab goes onto the caller's host grid (accumulate semantics, cu:542, 91-93): in place over PCIe if it is pinned
        s = rtd_add_device_to_host(ctx, ctx->groupGrid + slabOff, hostDoseInOut + slabOff, slabN, &modified);
    }
    if (s == RTD_OK) { cudaError_t e = cudaEventRecord(evAll, ctx->stream); if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream); if (e != cudaSuccess) s = fail(ctx, RTD_ERR_CUDA, cudaGetErrorString(e)); }
    if (s != RTD_OK) {
        const std::string msg = ctx->err;
        for (cudaStream_t bs : ctx->bstreams) cudaStreamSynchronize(bs);
        cudaStreamSynchronize(ctx->cstream); cudaStreamSynchronize(ctx->stream); cudaGetLastError();
        ctx->err = msg;
        return s;
    }
    if (ctx->groupFlag.as<int>()[0]) { ctx->groupFlag.as<int>()[0] = 0; return fail(ctx, RTD_ERR_CUDA, "rtd_plan_dose_group: a rank of the group did not reach the barrier within 10 s"); }
    s = doseFinish(ctx, nMine, tm, replayed ? evStart : evBegin, replayed ? evAll : evEnd, ct ? &io : nullptr, cs);
    if (tm) {
        float ms = 0; cudaEventElapsedTime(&ms, evStart, evAll); tm->msTotal = ms;      // zeroing, barriers, push and host tail included
        if (modified) { tm->h2dBytes += modified * 4; tm->d2hBytes += modified * 4; tm->doseVoxelsModified = modified; }
        tm->kernelLaunches += (world > 1 ? 2 : 0) + ((rank != 0 && slabN) ? 1 : 0) + ((rank == 0 && hostDoseInOut && slabN) ? 1 : 0);
    }
    return s;
}

rtd_status rtd_get_beam_report(const rtd_ctx* ctx, int beamNo, rtd_beam_report* out) {
    if (!ctx || !out || beamNo < 0 || beamNo >= ctx->nBeamsLast) return RTD_ERR_INVALID;
    *out = ctx->ws[beamNo]->rep;
    return RTD_OK;
}

rtd_status rtd_get_layer_report(const rtd_ctx* ctx, int beamNo, int layerNo, rtd_layer_report* out) {
    if (!ctx || !out || beamNo < 0 || beamNo >= ctx->nBeamsLast) return RTD_ERR_INVALID;
    const BeamWs& w = *ctx->ws[beamNo];
    if (layerNo < 0 || layerNo >= (int)w.lrep.size()) return RTD_ERR_INVALID;
    *out = w.lrep[layerNo];
    return RTD_OK;
}

int rtd_debug_ct_boxes(const rtd_beam* b, int nx, int ny, int nz, int* boxes6, int maxBoxes) {
    if (!b || !b->sigmas || b->nLayers <= 0 || b->sx <= 0 || b->sy <= 0 || nx <= 0 || ny <= 0 || nz <= 0 || b->steps == 0) return -1;
    rtd_opts o; rtd_default_opts(&o);
    float mx = b->sigmas[0], my = b->sigmas[1];
    for (int l = 1; l < b->nLayers; ++l) { mx = std::max(mx, b->sigmas[2 * l]); my = std::max(my, b->sigmas[2 * l + 1]); }
    const RayGrid grid = makeRayGrid(b->spotDelta, b->spotOffset, b->sx, b->sy, b->raySpacing, mx, my, o.convSigmaCutoff);
    if (grid.dimX <= 0 || grid.dimY <= 0) return -1;
    Affine gtii; gtii.m = m3(b->gantryToImIdxM); gtii.v = v3(b->gantryToImIdxV[0], b->gantryToImIdxV[1], b->gantryToImIdxV[2]);
    const TraceParams tp = makeTraceParams(1.0f, 1.0f, b->steps, grid.rayIdxToGantry, b->sourceDist, gtii);
    const int dims[3] = {nx, ny, nz};
    std::vector<Box> v;
    ctBoxesOfBeam(tp, grid.dimX, grid.dimY, (int)b->steps, dims, v);
    for (size_t i = 0; i < v.size() && (int)i < maxBoxes && boxes6; ++i)
        for (int d = 0; d < 3; ++d) { boxes6[6 * i + d] = v[i].lo[d]; boxes6[6 * i + 3 + d] = v[i].hi[d]; }
    return (int)v.size();
}

rtd_status rtd_debug_fetch(rtd_ctx* ctx, int beamNo, rtd_fetch_what what, int layerNo, void* dst, size_t dstBytes) {
    if (!ctx || !dst || beamNo < 0 || beamNo >= ctx->nBeamsLast) return RTD_ERR_INVALID;
    BeamWs& w = *ctx->ws[beamNo];
    const size_t R = w.R, S = w.S, nz = w.nz > 0 ? w.nz : 0;
    const size_t tiles = size_t(w.grid.dimX / 32) * (w.grid.dimY / 8);
    const void* src = nullptr; size_t bytes = 0;
    const bool perLayer = (what == RTD_FETCH_IDD || what == RTD_FETCH_RSIGMA || what == RTD_FETCH_FIRST_PASSIVE || what == RTD_FETCH_TILE_RADIUS);
    if (perLayer && (layerNo < w.l0 || layerNo >= w.l0 + w.Lc)) return fail(ctx, RTD_ERR_INVALID, "rtd_debug_fetch: bad layer");
    layerNo -= w.l0;
    if ((perLayer || what == RTD_FETCH_BEV_DOSE || what == RTD_FETCH_RAY_WEIGHTS) && !w.live) return fail(ctx, RTD_ERR_STATE, "rtd_debug_fetch: beam has no live steps");
    switch (what) {
        case RTD_FETCH_DENSITY: src = w.density.p; bytes = R * S * 4; break;
        case RTD_FETCH_WEPL: src = w.wepl.p; bytes = R * S * 4; break;
        case RTD_FETCH_FIRST_INSIDE: src = w.firstIn.p; bytes = R * 4; break;
        case RTD_FETCH_FIRST_OUTSIDE: src = w.firstOut.p; bytes = R * 4; break;
        case RTD_FETCH_WEPL_MIN: src = w.weplMin.as<float>() + 2; bytes = S * 4; break;
        case RTD_FETCH_RAY_WEIGHTS: src = w.rayW.p; bytes = size_t(w.Lc) * R * 4; break;
        case RTD_FETCH_IDD: src = w.idd.as<float>() + size_t(layerNo) * nz * R; bytes = nz * R * 4; break;
        case RTD_FETCH_RSIGMA: src = w.rsig.as<float>() + size_t(layerNo) * nz * R; bytes = nz * R * 4; break;
        case RTD_FETCH_FIRST_PASSIVE: src = w.firstPassive.as<int>() + size_t(layerNo) * R; bytes = R * 4; break;
        case RTD_FETCH_BEV_DOSE: src = w.bev.p; bytes = size_t(w.Zg) * (w.grid.dimY + 64) * (w.grid.dimX + 64) * 4; break;
        case RTD_FETCH_TILE_RADIUS: src = w.radMap.as<uint8_t>() + size_t(layerNo) * nz * tiles; bytes = nz * tiles; break;
        default: return fail(ctx, RTD_ERR_INVALID, "rtd_debug_fetch: unknown item");
    }
    if (dstBytes < bytes) return fail(ctx, RTD_ERR_INVALID, "rtd_debug_fetch: destination too small");
    RTD_CUDA(cudaSetDevice(ctx->device));
    RTD_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    return RTD_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------
// rtd_add_device_to_host_noalloc: the multi-GPU tail (rtd_add_device_to_host) for callers that must not allocate - code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc. No counterpart in the reference.
// Nothing on this path calls cudaMalloc / cudaFree / cudaMallocHost or any allocating library routine.
//
// The scratch of the kernel - one partial count per block, a ticket and the total - is a file-scope device array of fixed size, held
// by every GPU that loads the library until the process exits (33 024 bytes). RTD_NOALLOC_CHUNK voxels per block and
// RTD_NOALLOC_MAX_BLOCKS blocks give the stated limit RTD_NOALLOC_MAX_VOXELS = 2^28 voxels (1 GiB of fp32); larger n is refused.
// In host code the array's NAME is a host-side stand-in, not a device address: the address is taken once per device with
// cudaGetSymbolAddress (that device current) and passed to the kernel. Every call on a device shares the array and it keeps the last
// call's contents: the ticket and the total are cleared at the start of every call (the partial counts are overwritten before they are
// read); two calls that overlap on the GPU (different streams / contexts of one device) race on it and must be serialised by the caller.
namespace {

constexpr size_t RTD_NOALLOC_CHUNK = 65536;                  // voxels per block
constexpr size_t RTD_NOALLOC_MAX_BLOCKS = 4096;
static_assert(RTD_NOALLOC_MAX_VOXELS == RTD_NOALLOC_CHUNK * RTD_NOALLOC_MAX_BLOCKS, "include/rtdose.h states the limit");      // 2^28
constexpr int RTD_NOALLOC_HEAD = 32;                         // [0] ticket, [1] total, rest of the first 256 bytes unused
static __device__ __align__(256) unsigned long long g_noallocScratch[RTD_NOALLOC_HEAD + RTD_NOALLOC_MAX_BLOCKS];

// dst[i] += src[i] where src[i] != 0 (add_nonzero_kernel, kernels.cuh), block b owning voxels [b * CHUNK, (b + 1) * CHUNK); the count of
// voxels touched is reduced per block into scratch[HEAD + b] and summed by the block that arrives last (ticket), without float or
// 64-bit atomics on the caller's memory: the total is scratch[1].
__global__ void __launch_bounds__(256)
add_nonzero_counted_kernel(const float* __restrict__ src, float* __restrict__ dst, const size_t n, unsigned long long* __restrict__ scratch) {
    __shared__ unsigned sCount[8];
    __shared__ bool sLast;
    const size_t lo = (size_t)blockIdx.x * RTD_NOALLOC_CHUNK, hi = min(n, lo + RTD_NOALLOC_CHUNK);
    unsigned nMod = 0;
    for (size_t i = lo + threadIdx.x; i < hi; i += 256) {
        const float v = src[i];
        if (v != 0.0f) { dst[i] += v; ++nMod; }
    }
    nMod = __reduce_add_sync(0xffffffffu, nMod);
    if ((threadIdx.x & 31) == 0) sCount[threadIdx.x >> 5] = nMod;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned c = 0;
        for (int w = 0; w < 8; ++w) c += sCount[w];
        scratch[RTD_NOALLOC_HEAD + blockIdx.x] = c;
        __threadfence();
        sLast = atomicAdd(scratch, 1ull) == (unsigned long long)(gridDim.x - 1);
    }
    __syncthreads();
    if (!sLast) return;
    __threadfence();
    unsigned long long t = 0;
    for (unsigned b = threadIdx.x; b < gridDim.x; b += 256) t += *(volatile unsigned long long*)(scratch + RTD_NOALLOC_HEAD + b);
    t = __reduce_add_sync(0xffffffffu, (unsigned)t);        // <= 4096 blocks x 65536 voxels = 2^28: fits 32 bits
    __shared__ unsigned sTot[8];
    if ((threadIdx.x & 31) == 0) sTot[threadIdx.x >> 5] = (unsigned)t;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long tot = 0;
        for (int w = 0; w < 8; ++w) tot += sTot[w];
        scratch[1] = tot;
    }
}

// Device address of the scratch array on `device` (which must be current), resolved once per device.
cudaError_t noallocScratch(int device, unsigned long long** out) {
    constexpr int MAXDEV = 64;
    static unsigned long long* ptr[MAXDEV] = {};
    static std::mutex mu;
    if (device < 0 || device >= MAXDEV) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> g(mu);
    if (!ptr[device]) {
        unsigned long long* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noallocScratch);
        if (e != cudaSuccess) return e;
        ptr[device] = p;
    }
    *out = ptr[device];
    return cudaSuccess;
}

}  // namespace

extern "C" rtd_status rtd_add_device_to_host_noalloc(rtd_ctx* ctx, const float* dSrc, float* dst, size_t n, int64_t* modifiedOut) {
    if (!ctx) return RTD_ERR_INVALID;
    if (!dSrc || !dst) return fail(ctx, RTD_ERR_INVALID, "rtd_add_device_to_host_noalloc: null pointer");
    if (n > RTD_NOALLOC_MAX_VOXELS)
        return fail(ctx, RTD_ERR_INVALID, "rtd_add_device_to_host_noalloc: n exceeds the fixed scratch (at most 268435456 voxels = 2^28); use rtd_add_device_to_host");
    RTD_CUDA(cudaSetDevice(ctx->device));
    if (modifiedOut) *modifiedOut = 0;
    if (n == 0) return RTD_OK;
    // destination: device memory, or page-locked host memory the GPU can address (no staging buffer may be allocated here)
    cudaPointerAttributes a0, a1;
    float* d = nullptr;
    if (cudaPointerGetAttributes(&a0, dst) == cudaSuccess && cudaPointerGetAttributes(&a1, dst + (n - 1)) == cudaSuccess && a0.type == a1.type &&
        (a0.type == cudaMemoryTypeHost || a0.type == cudaMemoryTypeDevice || a0.type == cudaMemoryTypeManaged) && a0.devicePointer &&
        (char*)a1.devicePointer - (char*)a0.devicePointer == (ptrdiff_t)((n - 1) * 4))
        d = (float*)a0.devicePointer;
    cudaGetLastError();
    if (!d) return fail(ctx, RTD_ERR_INVALID, "rtd_add_device_to_host_noalloc: the destination must be device memory or page-locked host memory (a pageable grid needs a staging buffer: rtd_add_device_to_host)");
    unsigned long long* scratch = nullptr;
    RTD_CUDA(noallocScratch(ctx->device, &scratch));
    RTD_CUDA(cudaMemsetAsync(scratch, 0, 16, ctx->stream));                 // ticket and total start at zero whatever the last call left
    const unsigned blocks = (unsigned)((n + RTD_NOALLOC_CHUNK - 1) / RTD_NOALLOC_CHUNK);
    add_nonzero_counted_kernel<<<blocks, 256, 0, ctx->stream>>>(dSrc, d, n, scratch);
    RTD_CUDA(cudaGetLastError());
    unsigned long long total = 0;
    RTD_CUDA(cudaMemcpyAsync(&total, scratch + 1, 8, cudaMemcpyDeviceToHost, ctx->stream));
    RTD_CUDA(cudaStreamSynchronize(ctx->stream));
    if (modifiedOut) *modifiedOut = (int64_t)total;
    return RTD_OK;
}
