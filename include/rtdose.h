/*
 * rtdose.h - C ABI of the B200-native proton pencil-beam dose engine (librtdose.so).
 *
 * This is the drop-in boundary for the reference's single hot-path entry point
 *     void cudaWrapperProtons(HostPinnedImage3D<float>* imVol, HostPinnedImage3D<float>* doseVol,
 *                             const std::vector<BeamSettings> beams, const EnergyStruct iddData,
 *                             std::ostream& outStream);          (reference: src/kernel_wrapper.cuh:161,
 *                                                                 implemented src/kernel_wrapper.cu:381-1369)
 * Plain pointers and sizes only; no C++/torch types. include/rtdose_shim.hpp wraps these calls in a C++
 * function with the reference's exact signature and value types (BeamSettings, EnergyStruct,
 * HostPinnedImage3D), see INTEGRATION.md.
 *
 * Conventions (identical to the reference):
 *   - CT voxels are float(HU + 1000), x fastest, dims (nx,ny,nz)       (src/dicom_reader.cpp:104, src/main.cu:42)
 *   - dose is ACCUMULATED onto the caller's content                     (src/kernel_wrapper.cu:542, 91-93)
 *   - spot weights are [layer][sy][sx]                                   (src/beam_settings.h:21)
 *   - the beam travels along -z gantry (spotDelta[2] < 0)                (src/fill_idd_and_sigma_params.cu:22)
 * Ownership: every pointer is borrowed for the duration of the call; the library frees only what it
 * allocated and never resets the device (the reference deletes its inputs and calls cudaDeviceReset(),
 * src/kernel_wrapper.cu:856,1366-1368 - the C++ shim reproduces that for main.cu-style callers).
 * Threading: one context per device; calls on one context must not overlap.
 *
 * Limits the reference does not have (RTD_ERR_INVALID beyond them): at most 255 energy layers per beam (the superposition's layer list
 * packs the layer number into 8 bits), at most 4095 tracer steps per beam, at most 4095 tiles of 32 x 8 rays per beam (a 1024 x 1024-ray
 * grid is refused), spot grid at most 2^24 spots per layer. Every beam is validated before anything is launched: non-finite or
 * non-positive spot pitch, ray spacing, sigmas, non-finite energies / transforms, a zero source distance are RTD_ERR_INVALID.
 * Workspaces are sized for the worst case (every tracer step live) and only grow: 2 x 4 B x layers x steps x rays per beam dominate.
 * They come out of a few large chunks per context (one per plan size seen) and go back to the driver in rtd_destroy only; a buffer
 * that has to grow takes a new region 1.5 x as large, so what a context holds stays below about twice its largest plan's need.
 */
#ifndef RTDOSE_H
#define RTDOSE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTD_MAX_SUPERP_R 32   /* maxSuperpR, src/kernel_wrapper.cuh:26 */
#define RTD_TILE_X 32         /* superpTileX, src/kernel_wrapper.cuh:27 */
#define RTD_TILE_Y 8          /* superpTileY, src/kernel_wrapper.cuh:28 */
#define RTD_HIST_BINS 34      /* maxSuperpR + 2 radius counters, src/kernel_wrapper.cu:734 */

typedef struct rtd_ctx rtd_ctx;

typedef enum {
    RTD_OK = 0,
    RTD_ERR_INVALID = 1,   /* bad argument (replaces std::runtime_error("Empty list"), src/vector_find.h:22-25, etc.) */
    RTD_ERR_CUDA = 2,      /* CUDA runtime / launch error (replaces cudaErrchk throw, src/cuda_errchk.cu:12-22) */
    RTD_ERR_RADIUS = 3,    /* a tile needs a superposition radius > 32 (replaces throw("..."), src/kernel_wrapper.cu:965) */
    RTD_ERR_STATE = 4,     /* LUTs / CT / dose grid not set */
    RTD_ERR_NO_DEVICE = 5  /* no CUDA device: the product has no CPU fallback */
} rtd_status;

/* POD mirror of EnergyStruct (reference: src/energy_struct.h:13-38). */
typedef struct {
    int nEnergySamples;          /* depth samples per energy (1024) */
    int nEnergies;               /* 147 */
    const float* energiesPerU;   /* [nEnergies] ascending */
    const float* peakDepths;     /* [nEnergies] mm */
    const float* scaleFacts;     /* [nEnergies] depth-sample per mm WEPL */
    const float* ciddMatrix;     /* [nEnergies][nEnergySamples] cumulative integral depth dose */
    int nDensitySamples; float densityScaleFact; const float* densityVector;   /* HU+1000 -> g/cm3 */
    int nSpSamples;      float spScaleFact;      const float* spVector;        /* HU+1000 -> relative stopping power */
    int nRRlSamples;     float rRlScaleFact;     const float* rRlVector;       /* density*scale -> 1/X0 */
    /* NUCLEAR_CORR builds only (src/energy_struct.h:33-36, read by src/energy_reader.cpp:103-162): optional, NULL otherwise */
    const float* nucWeightMatrix;    /* [nEnergies][nEnergySamples] share of the dose attributed to the nuclear halo */
    const float* nucSqSigmaMatrix;   /* [nEnergies][nEnergySamples] variance of the halo (accepted, not used: see rtd_opts.nuclearCorr) */
} rtd_luts;

/* POD mirror of BeamSettings (reference: src/beam_settings.h:101-109; ctor :32). */
typedef struct {
    const float* weights;        /* [nLayers][sy][sx] spot weights (host) */
    int sx, sy, nLayers;
    const float* energies;       /* [nLayers] MeV/u */
    const float* sigmas;         /* [nLayers][2] spot sigma (x,y) at iso in air, mm */
    float raySpacing[2];         /* ray pitch at iso, mm */
    unsigned int steps;          /* tracer steps */
    float sourceDist[2];         /* apparent source-to-iso distance x,y (INFINITY = parallel) */
    float spotDelta[3];          /* spotIdxToGantry delta (z: signed step length, negative) */
    float spotOffset[3];         /* spotIdxToGantry offset (z: start plane) */
    float gantryToImIdxM[9];  float gantryToImIdxV[3];     /* row-major 3x3 + translation */
    float gantryToDoseIdxM[9]; float gantryToDoseIdxV[3];
    /* Layer shard: compute only energy layers [layerBegin, layerEnd) of this beam (0,0 = all). The ray grid, the
     * tracer and every cut-off still come from the WHOLE beam, so the doses of disjoint shards add up to the dose
     * of the whole beam (multi-GPU work items). Not part of the reference's BeamSettings. */
    int layerBegin, layerEnd;
} rtd_beam;

/* Run-time equivalents of the reference's compile-time flags (reference: CMakeLists.txt:36-79). */
typedef struct {
    float bpDepthCutoff;     /* BP_DEPTH_CUTOFF   1.05 */
    float convSigmaCutoff;   /* CONV_SIGMA_CUTOFF 3.0  */
    float ksSigmaCutoff;     /* KS_SIGMA_CUTOFF   3.0  */
    float rayWeightCutoff;   /* RAY_WEIGHT_CUTOFF 1.0  */
    int   minTilesInBatch;   /* 16, src/kernel_wrapper.cuh:29 */
    int   ks2BodyStep;       /* measurement hook of the superposition kernels (csrc/ks2.cuh, Ks2Params::bodyStep; results unchanged):
                                0 (default) = chosen per beam on the device, 1 | 2 | 4 = a radius <= 16 runs in the unrolled body of the
                                next multiple */
    int   serialBeams;       /* 1: run the beams one after the other on one stream (stage timings then do not overlap);
                                0 (default): one stream per beam, transfers into the dose grid serialised in beam order */
    int   noZeroCopyDose;    /* 0 (default): a PINNED host dose grid is accumulated in place by the transfer kernels (only voxels
                                that receive dose cross PCIe); 1: always stage dose sub-volumes through device memory */
    int   noNozzle;          /* 1: the reference's -DNO_NOZZLE build: no in-air spot-size growth (fill_idd_and_sigma_params.cu:76-82) */
    int   doseToMedium;      /* 1: the reference built with DOSE_TO_WATER off: voxel mass = density x volume (kernel_wrapper.cu:314-318) */
    int   atomicDose;        /* 1: the transfers add with red.global.add.f32: the dose grid may then be ONE buffer shared by all
                                ranks of a node (rtd_shared_dose_*), written over NVLink - no separate reduce collective */
    int   ctPull;            /* rtd_plan_dose, 1: a PINNED host CT is read by the GPU itself (ct_pull_rows_kernel): per CT row only the x
                                interval the tracers can touch crosses PCIe (C3: 51 MB instead of 60 MB of boxes). Measured SLOWER end to
                                end on B200 (3.2 vs 2.9 ms: the pull kernels compete with the dose kernels for SMs, the copy engines do
                                not), so the default 0 keeps the axis-aligned sub-volume copies */
    int   nuclearCorr;       /* the reference's NUCLEAR_CORR build option (CMakeLists.txt:57-79) as shipped: 0 = off (its default), 1 = SOUKUP,
                                2 = FLUKA, 3 = GAUSS_FIT (the reference numbers them 0 / 1 / 2). Needs rtd_luts.nucWeightMatrix of that model.
                                What the option does in the reference: E_s and the empirical widening of the depth scan change per model
                                (kernel_wrapper.cu:232-240), GAUSS_FIT drops the post-peak sigma correction (cu:300-302) and shrinks the entry
                                sigmas by 3 % (cu:842-847), and the primary dose is scaled by 1 - nucWeight (cu:320-341). Its halo pass is
                                constructed with nucMemoryStep = 0 (cu:925), so every depth step of a halo pencil beam lands in BEV slice 0 and
                                the second superposition / nucTransfDiv (cu:978-996, 1058-1092, 1220-1255) deposit nothing unless the beam's
                                first live step is step 0; that one case (beamFirstInside == 0) is rejected with RTD_ERR_INVALID */
    int   noGraph;           /* 0 (default): a device-resident plan with several beams (rtd_dose_device) is captured once into a CUDA graph and
                                replayed while the plan's geometry, options, buffers and LUTs stay the same (spot weights may change: they are
                                copied from a staging buffer at every replay); stage times of rtd_timing are 0 for replayed plans. 1: always
                                issue the launches one by one */
    int   fullIntermediates; /* 1: the depth scan writes idd / rSigmaEff of EVERY (ray, step) as the reference does, for rtd_debug_fetch
                                (RTD_FETCH_IDD / RTD_FETCH_RSIGMA). 0 (default): 32-ray row segments that hold no dose at a step are
                                not written (the superposition never reads them; ~60 % of the scan's HBM writes on a head plan) and
                                fetch as unspecified values */
    int   tracerBrick;       /* tracer variant: 1 = the CT is staged per (32 x 8 rays, NS steps) brick in shared memory by TMA
                                from a linear fp32 copy and filtered by hand with the texture unit's integer weights - bit-identical to the
                                texture path for integer-valued CT (HU + 1000 of a DICOM series); a beam whose bricks hold a non-integer
                                texel is recomputed through the texture unit by a gated second launch. 0 = texture fetches only.
                                -1 (what rtd_default_opts sets) = the library's choice */
} rtd_opts;

/* Stage timing of the last rtd_dose* call (CUDA events; with concurrent beams the per-stage figures are sums over the
 * beams' streams and overlap in time, so they can exceed msTotal), the counterpart of the
 * reference's FINE_GRAINED_TIMING report (src/kernel_wrapper.cu:1291-1311). */
typedef struct {
    float msTotal;       /* first event -> last event of the call */
    float msUpload;      /* host->device copies inside the call (CT if dirty, dose-in, weights) */
    float msTrace;       /* K1 trace + reductions, all beams */
    float msSetup;       /* host logic + per-layer parameter upload + BEV clear */
    float msConv;        /* K3a spot -> ray weights */
    float msIddSigma;    /* K2 */
    float msRadius;      /* tile radius + batching */
    float msSuperp;      /* K3b kernel superposition */
    float msTransfer;    /* K4 BEV -> dose grid */
    float msDownload;    /* device->host dose */
    int64_t updates;     /* U = sum_beams sum_layers Rx*Ry*(layerFirstPassive - beamFirstInside) */
    int64_t ksTaps;      /* sum over updates of (2*rb+1)^2, rb = batch radius */
    int     kernelLaunches;
    int     hostSyncs;
    int64_t h2dBytes;    /* bytes copied host->device inside the call (CT sub-volumes, dose-in sub-volumes, weights, parameters) */
    int64_t d2hBytes;    /* bytes copied device->host inside the call (dose sub-volumes, reductions, reports) */
    int64_t doseVoxelsModified;   /* dose-grid voxels that received dose (summed over beams) */
    int     ksLaunchesWithWork;   /* superposition launches (beam x radius class) that processed at least one item */
    int     tracerBrickBeams;     /* beams traced by the TMA-brick tracer whose result stood (no fall-back to the texture path) */
    int     tracerOutsideSamples; /* samples of those beams that missed their brick and were filtered from global memory (expected: 0) */
    int     reserved0;
} rtd_timing;

typedef struct {
    int lSteps, bSteps, rSteps, tSteps;      /* src/kernel_wrapper.cu:650-653 */
    int rayDimX, rayDimY, steps, nLayers;    /* src/kernel_wrapper.cu:659 */
    int beamFirstInside, beamFirstOutside;   /* src/kernel_wrapper.cu:781-787 */
    int beamFirstGuaranteedPassive;          /* src/kernel_wrapper.cu:796 */
    int beamFirstCalculatedPassive;          /* src/kernel_wrapper.cu:955-957 */
    int bboxMin[3], bboxMax[3];              /* src/kernel_wrapper.cu:1207-1208 */
    float entryZ;
    int64_t updates;
} rtd_beam_report;

typedef struct {
    float energyIdx, energyScaleFact, peakDepth;   /* src/kernel_wrapper.cu:835-837 */
    float entrySigmaX, entrySigmaY;                /* src/kernel_wrapper.cu:840-841 */
    int afterLastStep;                             /* src/kernel_wrapper.cu:923-924 */
    int layerFirstPassive;                         /* src/kernel_wrapper.cu:952-954 */
    int radiusHist[RTD_HIST_BINS];                 /* tilePrimRadCtrs, src/kernel_wrapper.cu:959-963 */
    int batchRadius[RTD_HIST_BINS];                /* radius actually applied to tiles of own radius r (cu:966-976) */
} rtd_layer_report;

typedef enum {
    RTD_FETCH_DENSITY = 0,       /* float [steps][Ry][Rx]                         (bevDensity) */
    RTD_FETCH_WEPL = 1,          /* float [steps][Ry][Rx]                         (bevCumulSp) */
    RTD_FETCH_FIRST_INSIDE = 2,  /* int   [Ry][Rx] */
    RTD_FETCH_FIRST_OUTSIDE = 3, /* int   [Ry][Rx] */
    RTD_FETCH_WEPL_MIN = 4,      /* float [steps] */
    RTD_FETCH_RAY_WEIGHTS = 5,   /* float [L][Ry][Rx] */
    RTD_FETCH_IDD = 6,           /* float [nz][Ry][Rx] of one layer, nz = Zg - beamFirstInside, slice 0 = step beamFirstInside */
    RTD_FETCH_RSIGMA = 7,        /* float [nz][Ry][Rx] of one layer */
    RTD_FETCH_FIRST_PASSIVE = 8, /* int   [Ry][Rx] of one layer */
    RTD_FETCH_BEV_DOSE = 9,      /* float [Zg][Ry+64][Rx+64] */
    RTD_FETCH_TILE_RADIUS = 10   /* uint8 [nz][Ry/8][Rx/32] of one layer (own radius, before batching; 254 = live tile without dose, 255 = not live) */
} rtd_fetch_what;

void        rtd_default_opts(rtd_opts* o);
rtd_status  rtd_create(rtd_ctx** out, int device);    /* RTD_ERR_NO_DEVICE without a CUDA device (no CPU fallback) */
/* Same, with the caller's stream (a cudaStream_t of `device`) as the context's main stream: the caller keeps owning it (rtd_destroy
 * leaves it alone) and can order its own work - NCCL, torch, copies - on it around the rtd_* calls without a host sync. */
rtd_status  rtd_create_on_stream(rtd_ctx** out, int device, void* cudaStream);
void        rtd_destroy(rtd_ctx* ctx);
const char* rtd_last_error(const rtd_ctx* ctx);   /* message of the last non-OK status ("" if none) */
const char* rtd_version(void);

/* Uploads the LUTs (textures: IDD 2-D clamp, density/SP/1/X0 1-D clamp; src/kernel_wrapper.cu:453-537). */
rtd_status rtd_set_luts(rtd_ctx* ctx, const rtd_luts* luts);
/* Uploads the CT (3-D texture, linear, border; src/kernel_wrapper.cu:420-451). huPlus1000 is a HOST pointer. */
rtd_status rtd_set_ct(rtd_ctx* ctx, const float* huPlus1000, int nx, int ny, int nz);
/* Same, from a DEVICE pointer (device-resident benchmarking / multi-GPU broadcast). */
rtd_status rtd_set_ct_device(rtd_ctx* ctx, const float* dHuPlus1000, int nx, int ny, int nz);
/* Declares the dose grid dims (the reference takes them from doseVol->getDims()). */
rtd_status rtd_set_dose_grid(rtd_ctx* ctx, int nx, int ny, int nz);

/* The dose call. doseInOut: HOST float[nz][ny][nx], accumulated in place (H2D + all beams + D2H inside). */
rtd_status rtd_dose(rtd_ctx* ctx, const rtd_beam* beams, int nBeams, float* doseInOut,
                    const rtd_opts* opts, rtd_timing* timing);
/* Same with a DEVICE dose buffer (accumulated in place, no host copies of the dose); the call returns after the
 * context's stream has been synchronised. */
rtd_status rtd_dose_device(rtd_ctx* ctx, const rtd_beam* beams, int nBeams, float* dDoseInOut,
                           const rtd_opts* opts, rtd_timing* timing);

/* The one-call equivalent of cudaWrapperProtons (src/kernel_wrapper.cu:381-1369): HOST CT + HOST dose, accumulated in
 * place. Where the reference uploads the whole CT (cu:420-451) and the whole dose grid (cu:542) and downloads the whole
 * dose grid (cu:1357), this call moves only the sub-volumes the plan can touch: the CT boxes that enclose every
 * tracer sample of each beam (the fan frustum cut into depth segments) before the tracers, and the dose boxes that
 * enclose each beam's haloed BEV volume (kernel_wrapper.cu:1185-1210) - uploaded while the beams compute, downloaded
 * after the last transfer. Voxels outside those boxes are never read or written by any kernel, so the result equals
 * rtd_set_ct + rtd_dose. After this call the context's CT is only PARTIALLY resident: rtd_dose / rtd_dose_device
 * return RTD_ERR_STATE until rtd_set_ct* is called again. Pinned (cudaHostRegister'ed, like HostPinnedImage3D,
 * src/host_image_3d.cuh:24) host buffers make the copies asynchronous. Dose dims come from rtd_set_dose_grid. */
rtd_status rtd_plan_dose(rtd_ctx* ctx, const float* ctHuPlus1000, int nx, int ny, int nz, const rtd_beam* beams, int nBeams,
                         float* doseInOut, const rtd_opts* opts, rtd_timing* timing);
/* Same CT handling (HOST CT, sub-volumes only) with a DEVICE dose buffer: a rank of a multi-GPU plan computes its work
 * items into its partial dose grid, which is then reduced over NCCL by the caller. */
rtd_status rtd_plan_dose_device(rtd_ctx* ctx, const float* ctHuPlus1000, int nx, int ny, int nz, const rtd_beam* beams, int nBeams,
                                float* dDoseInOut, const rtd_opts* opts, rtd_timing* timing);

/* Multi-GPU without a separate collective ("NVLS / peer stores issued from the transfer kernel"): one rank
 * creates the plan's dose grid and exports a 64-byte CUDA IPC handle, the other ranks of the node map it, and every rank
 * runs rtd_dose_device / rtd_plan_dose_device on the mapped pointer with opts.atomicDose = 1: the transfer kernels of all
 * GPUs accumulate into the one grid with float reductions over NVLink / NVSwitch while they compute. The caller orders
 * "grid zeroed" before and "all ranks done" after (a stream-ordered barrier). close: owner = 1 frees, 0 unmaps. */
rtd_status rtd_shared_dose_create(rtd_ctx* ctx, size_t nFloats, float** dPtr, unsigned char* handle64);
rtd_status rtd_shared_dose_open(rtd_ctx* ctx, const unsigned char* handle64, float** dPtr);
rtd_status rtd_shared_dose_close(rtd_ctx* ctx, float* dPtr, int owner);
/* Sparse push of a rank's partial grid onto the shared one (the faster of the two peer variants): adds the
 * float4 groups of dSrc[0..n) that hold dose onto dPeerDst[0..n) with 16-byte vector reductions over NVLink. Asynchronous
 * on rtd_stream; several ranks may push concurrently. Both pointers 16-byte aligned. */
rtd_status rtd_push_dose_to_peer(rtd_ctx* ctx, const float* dSrc, float* dPeerDst, size_t n);

/* ---- One plan on the GPUs of a node, one process / context per GPU. The reference's beam loop (src/kernel_wrapper.cu:601)
 * and layer loop (cu:916) are serial on one GPU; dose is linear in both, so (beam, layer range) work items are independent until the sum.
 *
 * rtd_partition / rtd_dose_slab are pure host functions (no device needed): the assignment of every (beam, layer) to a rank as
 * consecutive runs of near-equal cost (chunksPerBeam = 0; > 0 splits every beam into that many ranges, LPT-assigned), and the z-slab
 * of the dose grid the plan can reach. rtd_partition returns the number of work items (only maxItems are written), -1 on bad input;
 * rtd_dose_slab 1 / 0 (empty) / -1.
 *
 * rtd_group_create: rank 0 allocates the plan's ONE dose grid (nDoseFloats floats + a control block) and writes its 64-byte CUDA IPC
 * handle to handle64; the caller carries the handle to the other ranks (MPI, torch.distributed, a pipe ...), whose rtd_group_create
 * reads it and maps the grid over NVLink / NVSwitch peer memory. rtd_group_dose: the grid as seen from this rank.
 * rtd_group_barrier: a stream-ordered barrier of the ranks' streams (one remote atomic + polling in a one-thread kernel, no NCCL, no host
 * sync); every rank must call it the same number of times.
 * rtd_plan_dose_group: every rank passes the WHOLE plan. The call partitions it, computes this rank's work items (CT: the HOST volume
 * ctHuPlus1000 - only the sub-volumes of this rank's items are uploaded - or, if NULL, the CT resident from rtd_set_ct), sums the
 * partial doses in the group's grid (rank 0 computes into it directly, the others push the 16-byte groups that hold dose with vector
 * reductions over NVLink; two barriers order "grid zeroed" and "all pushed") and, on rank 0 with hostDoseInOut != NULL, adds the plan's
 * dose onto the caller's host grid. Returns when this rank's part is complete; on rank 0 the group's grid then holds the plan's dose
 * (zero outside rtd_dose_slab's slab is NOT guaranteed: only the slab is zeroed and summed). timing: this rank's own. */
typedef struct { int rank, beam, layerBegin, layerEnd; } rtd_work_item;
int        rtd_partition(const rtd_beam* beams, int nBeams, const rtd_luts* luts, int world, int chunksPerBeam, rtd_work_item* items, int maxItems);
int        rtd_dose_slab(const rtd_beam* beams, int nBeams, int nx, int ny, int nz, float convSigmaCutoff, int* z0, int* z1);
rtd_status rtd_group_create(rtd_ctx* ctx, int rank, int world, size_t nDoseFloats, unsigned char* handle64);
rtd_status rtd_group_destroy(rtd_ctx* ctx);
/* CT relay of the group (optional; new, the reference is single-GPU). Without it every rank of rtd_plan_dose_group uploads the CT
 * sub-volumes of every beam it holds a layer of over its own PCIe link (C3 on 8 GPUs: 168 MB for a plan whose sub-volumes unite to
 * 61 MB). rtd_group_ct_create gives this rank a staging buffer of `bytes` (the SAME size on every rank; a plan whose share per rank
 * exceeds it falls back to per-rank uploads) and writes its 64-byte CUDA IPC handle; after the caller has gathered the handles of all
 * ranks, in rank order, rtd_group_ct_open(handles = world x 64 bytes) maps the peers' buffers. From then on rtd_plan_dose_group with a
 * HOST CT deals the plan's boxes to the ranks by bytes: every rank uploads its share once, one stream-ordered group barrier later every
 * rank copies the boxes its beams touch from the owners' buffers into its CT array over NVLink. Freed by rtd_group_destroy. */
rtd_status rtd_group_ct_create(rtd_ctx* ctx, size_t bytes, unsigned char* handle64);
rtd_status rtd_group_ct_open(rtd_ctx* ctx, const unsigned char* handles);
float*     rtd_group_dose(rtd_ctx* ctx);
rtd_status rtd_group_barrier(rtd_ctx* ctx);
rtd_status rtd_plan_dose_group(rtd_ctx* ctx, const float* ctHuPlus1000, int nx, int ny, int nz, const rtd_beam* beams, int nBeams,
                               float* hostDoseInOut, const rtd_opts* opts, rtd_timing* timing);

/* Multi-GPU tail: after the ranks' partial dose grids have been summed on one device (NCCL reduce, issued
 * by the caller on rtd_stream), add the non-zero voxels of dSrc[0..n) onto the caller's HOST dose grid hostDst[0..n)
 * (the accumulate semantics of kernel_wrapper.cu:91-93 for the whole plan). A pinned hostDst is updated in place over
 * PCIe (only voxels that received dose travel); a pageable one is staged. modifiedOut (optional): voxels updated. */
rtd_status rtd_add_device_to_host(rtd_ctx* ctx, const float* dSrc, float* hostDst, size_t n, int64_t* modifiedOut);
/* The same tail for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc). No counterpart in the reference. The call allocates nothing (no cudaMalloc,
 * cudaFree, cudaMallocHost or allocating library routine): the kernel's scratch (per-block counts, 33 024 bytes) is a fixed device array
 * inside the library, held by every GPU that loads it until the process exits. Limits: n <= RTD_NOALLOC_MAX_VOXELS (2^28 voxels = 1 GiB
 * of fp32), RTD_ERR_INVALID beyond; dst must be DEVICE memory or PAGE-LOCKED host memory (a pageable grid would need a staging buffer),
 * RTD_ERR_INVALID otherwise. Every call on a device shares the scratch, which keeps the last call's contents (what must start at zero
 * is cleared by the call itself): calls that overlap on the GPU - two contexts or streams of one device - race and must be serialised
 * by the caller. Same result as rtd_add_device_to_host bit for bit; returns after the stream has drained. */
#define RTD_NOALLOC_MAX_VOXELS ((size_t)1 << 28)
rtd_status rtd_add_device_to_host_noalloc(rtd_ctx* ctx, const float* dSrc, float* dst, size_t n, int64_t* modifiedOut);

/* Integer/index outputs of the last call (the reference's bit-exact quantities). */
rtd_status rtd_get_beam_report(const rtd_ctx* ctx, int beamNo, rtd_beam_report* out);
rtd_status rtd_get_layer_report(const rtd_ctx* ctx, int beamNo, int layerNo, rtd_layer_report* out);
/* Copies an intermediate of the last call to host memory (every beam keeps its arrays until the next call). */
rtd_status rtd_debug_fetch(rtd_ctx* ctx, int beamNo, rtd_fetch_what what, int layerNo, void* dst, size_t dstBytes);
/* Host geometry only (no device needed): the CT sub-volumes rtd_plan_dose would upload for this beam into a CT of
 * nx x ny x nz voxels, as inclusive boxes (x0, y0, z0, x1, y1, z1). Returns the number of boxes (may exceed maxBoxes; only
 * maxBoxes are written) or -1 on bad input. Lets CPU tests check that every texel the tracer can fetch is covered. */
int rtd_debug_ct_boxes(const rtd_beam* beam, int nx, int ny, int nz, int* boxes6, int maxBoxes);

/* The CUDA stream (cudaStream_t) all work of this context is issued on, for callers that order NCCL /
 * torch work after rtd_dose_device. */
void* rtd_stream(rtd_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* RTDOSE_H */
