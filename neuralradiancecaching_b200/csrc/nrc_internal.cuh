// nrc_internal.cuh -- shared declarations of the sm_100a NRC library (not part of the C-ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>
#include "../../include/nrc_b200.h"
#include "../../include/nrc_math.h"

// ---- error plumbing -------------------------------------------------------------------------------
void nrc_set_error(const char* fmt, ...);

#define NRC_CUDA(call)                                                                                    \
    do {                                                                                                  \
        cudaError_t e__ = (call);                                                                         \
        if (e__ != cudaSuccess) {                                                                         \
            nrc_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__));         \
            return (e__ == cudaErrorMemoryAllocation) ? NRC_ERR_OOM : NRC_ERR_CUDA;                        \
        }                                                                                                 \
    } while (0)

#define NRC_TRY(call)                           \
    do {                                        \
        nrc_status s__ = (call);                \
        if (s__ != NRC_OK) return s__;          \
    } while (0)

#define NRC_REQUIRE(cond, msg)                                     \
    do {                                                           \
        if (!(cond)) {                                             \
            nrc_set_error("%s:%d: %s", __FILE__, __LINE__, msg);   \
            return NRC_ERR_INVALID;                                \
        }                                                          \
    } while (0)

// ---- MLP arena layout (same packing as the C-ABI comment) -----------------------------------------
#define NRC_W_OFF(l) ((l) * 4096)
#define NRC_B_OFF(l) (20672 + (l) * 64)

// ---- kernel ids: launch counting and per-kernel CUDA-event timing (nrc_ctx_profile_*) ---------------
enum NrcKernelId {
    K_ENCODE_POS = 0, K_ENCODE_DIR, K_INFER_FP32, K_TRAIN_FP32, K_REDUCE_PARTIALS, K_TRANSPOSE_WEIGHTS, K_NRC_STEP, K_MRHE_STEP,
    K_INIT_NORMAL, K_RENDER_WALK, K_TRAIN_WALK, K_LIGHT_RENDER, K_LIGHT_TRAIN, K_RENDER_COMBINE, K_TRAIN_COMBINE, K_COMPOSITE,
    K_BUILD_BRICKS, K_VOLUME_SAMPLE, K_PACK_RECORDS, K_CDF_ROWS, K_CDF_Y, K_INFER_TC, K_TRAIN_TC, K_PACK_BF16, K_MISC, K_DP_REDUCE_STEP,
    K_ENCODE_TILES, K_MLP_STREAM, K_RENDER_WALK_TAIL, K_REDUCE_STEP,
    K_COUNT
};

struct NrcProfRec { int kid; cudaEvent_t a, b; };
// The train trace (walk -> light -> combine) reads no weights, so consecutive frames' traces are independent of each other: up to
// NRC_TRAIN_WALK_STREAMS train walks are in flight at once (a walk is ~300 warps of pure dependent latency, so several of them take
// barely longer than one), each filling its own buffer set; a set is reused once the train kernel that read it has finished.
#define NRC_RENDER_SETS 3          // buffer sets of the render trace: walk(k+2) beside light(k+1) beside inference / composite(k)
#define NRC_TRAIN_SETS 4
#define NRC_TRAIN_WALK_STREAMS 3
nrc_status nrc_renderer_resolve_pending(nrc_renderer* r);   // renderer.cu

// ---- handles ---------------------------------------------------------------------------------------
struct nrc_ctx {
    int device = 0;
    int smCount = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;       // second stream: render walk overlaps the training chain
    cudaStream_t stream3 = nullptr;       // third stream: the train TRACE of frame k+1 (needs no weights) runs ahead of frame k's MLP tail
    cudaStream_t streamCopy = nullptr;    // asynchronous read-back of finished frames
    cudaEvent_t evJoin = nullptr;
    cudaStream_t stream5 = nullptr;       // fifth stream: lights + combine of the render trace, beside the next frame's render walk (stream2)
    cudaEvent_t evRenderWalk = nullptr;
    cudaStream_t stream4 = nullptr;       // fourth stream: lights + combine of the train trace, beside the next frame's walk
    cudaEvent_t evTrainTrace = nullptr, evTrainWalk = nullptr, evTrainFree[NRC_TRAIN_SETS] = {nullptr, nullptr, nullptr, nullptr};
    cudaStream_t streamWalk[NRC_TRAIN_WALK_STREAMS] = {nullptr, nullptr, nullptr};   // [0] = stream3: consecutive frames' train walks on alternating streams
    nrc_mlp_precision precision = NRC_MLP_FP32;
    nrc_infer_mode inferMode = NRC_INFER_FUSED;   // nrc_ctx_create reads NRC_INFER_SPLIT
    uint64_t launches = 0;
    std::vector<nrc_renderer*> renderers;   // nrc_sync settles their deferred vertex-pool checks (renderer.cu)
    std::vector<const volatile unsigned*> dpFailWords;   // pinned time-out words of this context's nrc_dp objects (dp.cu): non-zero = a fused exchange failed closed
    // per-kernel event timing (off by default; bench.py turns it on over its timed region)
    bool profiling = false;
    std::vector<NrcProfRec> profOpen;       // recorded, not yet read
    std::vector<NrcProfRec> profFree;       // event pairs for reuse
    double profMs[K_COUNT] = {0};
    uint64_t profCount[K_COUNT] = {0};
    // pinned staging for [h] pointers
    void* pinned = nullptr;
    size_t pinnedBytes = 0;
    // device scratch for staged inputs/outputs of the stand-alone ops
    void* scratch[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    size_t scratchBytes[6] = {0, 0, 0, 0, 0, 0};
};

struct nrc_mlp {
    nrc_ctx* ctx = nullptr;
    float lr = 0.001f, wd = 0.0f, beta1 = 0.5f;
    float* params = nullptr;     // 20995 (+pad)
    float* grad = nullptr;       // 20995 negative-gradient accumulators + 1 mseLoss  (20996, padded)
    float* momentum = nullptr;   // 20995
    float* paramsT = nullptr;    // transposed weight copy Wt[l][in][out] for the forward GEMMs (5*4096 + 64*4)
    float* partial = nullptr;    // per-block partial gradients of the train kernel [maxBlocks][NRC_PARTIAL_STRIDE]
    int partialBlocks = 0;
    float* gScratch = nullptr;   // [rows][32] dL/dx0 of the hash features, phase 1 -> phase 2 of the train kernels (mrhe_scatter.cuh)
    size_t gScratchRows = 0;
    float* histPartial = nullptr;   // [smCount][T] per-CTA hash-gradient histograms of the last train launch (parity preset, small T)
    size_t histPartialFloats = 0;
    uint16_t* paramsBf16 = nullptr;  // bf16 operand images for the tcgen05 path (built on demand)
    // opt-in Adam: second moment + bias correction; the reference's step is first-moment momentum only (A16)
    bool adam = false;
    float beta2 = 0.999f, eps = 1e-8f;
    float* second = nullptr;     // 20995 (allocated when Adam is switched on)
    uint32_t steps = 0;
    void* featTiles = nullptr;       // bf16 operand-image tiles of the split inference path (grown on demand)
    size_t featBytes = 0;
    bool lossTotalSlot = false;  // a fused data-parallel exchange publishes the global loss in grad[20996] (dp.cu)
    bool dirtyT = true;          // paramsT needs a rebuild after params changed
    bool dirtyBf16 = true;       // paramsBf16 needs a rebuild
};

#define NRC_PARTIAL_STRIDE 21000   // 20995 grads + loss + pad

struct nrc_mrhe {
    nrc_ctx* ctx = nullptr;
    float lr = 0.01f, wd = 0.0f;
    uint32_t L = 16, T = 16384, F = 2, minRes = 16, maxRes = 512;
    uint32_t res[16] = {0};
    float* tables = nullptr;     // [L][T][F]
    float* delta = nullptr;      // [L][T][F]
    size_t count = 0;            // L*T*F
    // opt-in "fixed" preset: grid-space lerp factors on their own axes, per-feature errors, per-level gradient tables
    bool fixed = false;
    bool adam = false;
    float beta1 = 0.9f, beta2 = 0.999f, eps = 1e-8f;
    float* m1 = nullptr;         // first / second moments, L*T*F each (allocated when Adam is switched on)
    float* m2 = nullptr;
    uint32_t steps = 0;
    // mrhe-step.comp visits every table entry every frame (2 GiB of traffic at T = 2^22), but in the shipped arithmetic gradients only ever
    // land in the head [0, 2T).  With weight decay 0 an entry whose delta is 0 is left bit-identical by the step
    // (w + lr * (0 - w * 0) = w) once it has been scrubbed (bad(w) -> 0) -- so after ONE full step the tail is a fixed point and later
    // steps visit the head only.  Anything that can change the tail or the rule (set_tables, set_config, fixed preset, Adam) clears this.
    bool tailClean = false;
};

struct nrc_volume {
    nrc_ctx* ctx = nullptr;
    uint32_t nx = 0, ny = 0, nz = 0;
    // 2x2x2 corner bricks: cell (i,j,k), i in [-1,nx-1], holds the 8 UNORM8 texels it interpolates,
    // border texels are 0 (CLAMP_TO_BORDER, opaque black): one 8-byte load per density tap.
    uint2* bricks = nullptr;     // [(nz+1)][(ny+1)][(nx+1)]
    nrc_volume_params params;
};

struct nrc_envmap {
    nrc_ctx* ctx = nullptr;
    uint32_t w = 0, h = 0;
    float4* rgba = nullptr;
    float* invCdfX = nullptr;
    float* invCdfY = nullptr;
    float directStrength = 1.0f, hpmStrength = 8.0f;
    bool cdfHalf = false;        // opt-in: half of the env gather by inverse-CDF importance sampling (fwd.frag:727-744)
};

// Everything the trace kernels need, passed by value as a __grid_constant__ kernel parameter.
struct TraceParams {
    float invProjView[16];
    float camPos[3];
    float random[4];
    uint32_t useNN, showNonNN;
    float densityFactor, g;
    int noNnSpp, withNnSpp;
    float dirDir[3], dirStrength;
    float ptPos[3], ptStrength, ptColor[3];
    float envDirect, envHpm;
    float skySize[3];
    float exitJump;               // length(2*skySize), find_entry_exit
    int nx, ny, nz;
    const uint2* bricks;
    const float4* env;
    int envW, envH;
    const float* envInvCdfX;      // [envH][envW], non-null only when the CDF half of the env gather is switched on
    const float* envInvCdfY;      // [envH]
    uint32_t width, height;       // frame
    uint32_t trainW, trainH;
    float widthFactor, heightFactor;
};

// MRHE view handed to kernels by value.
struct MrheDev {
    const float* tables;
    float* delta;
    uint32_t L, T;
    uint32_t fixed;               // 0 = the reference's arithmetic, 1 = the opt-in fixed preset
    uint32_t res[16];
    float skySize[3];
};
// true the first time it is called for ctx's device with this flag array (one-time per-device setup such as cudaFuncSetAttribute)
inline bool first_on_device(const nrc_ctx* ctx, bool (&done)[64])
{
    const int dev = ctx->device;
    if (dev < 0 || dev >= 64) return true;
    if (done[dev]) return false;
    done[dev] = true;
    return true;
}

MrheDev make_mrhe_dev(const nrc_mrhe* m);

// ---- helpers implemented in capi.cu ---------------------------------------------------------------
void nrc_prof_begin(nrc_ctx* ctx, cudaStream_t st, int kid);
void nrc_prof_end(nrc_ctx* ctx, cudaStream_t st);
// Every kernel launch of the library goes through this: counts it, brackets it with CUDA events on the launching
// stream when profiling is on, and checks the launch.
#define NRC_LAUNCH(ctx, st, kid, ...)            \
    do {                                         \
        nrc_prof_begin((ctx), (st), (kid));      \
        __VA_ARGS__;                             \
        nrc_prof_end((ctx), (st));               \
        (ctx)->launches++;                       \
        NRC_CUDA(cudaGetLastError());            \
    } while (0)
bool nrc_is_device_ptr(const void* p);
// NRC_ERR_CUDA (with a message) once a fused data-parallel exchange of this context has given up on a peer (dp.cu): the replicas can
// no longer be trusted to be identical
nrc_status nrc_ctx_check_dp(nrc_ctx* ctx);
nrc_status nrc_scratch(nrc_ctx* ctx, int slot, size_t bytes, void** out);
// returns a device pointer holding the data of `p` (p itself when it already is one)
nrc_status nrc_stage_in(nrc_ctx* ctx, int slot, const void* p, size_t bytes, const void** dev);
// returns a device pointer to write results to; nrc_stage_out_finish copies back when dst is host memory
nrc_status nrc_stage_out(nrc_ctx* ctx, int slot, void* dst, size_t bytes, void** dev);
nrc_status nrc_stage_out_finish(nrc_ctx* ctx, void* dst, const void* dev, size_t bytes);

// ---- kernels' host launchers (one per .cu) --------------------------------------------------------
// encode.cu
nrc_status launch_encode_pos(nrc_mrhe* m, const float* pos, size_t n, float* feat, uint32_t* idx);
nrc_status launch_encode_dir(nrc_ctx* ctx, const float* dir, size_t n, float* feat);
// mlp_fp32.cu
nrc_status launch_mlp_prepare(nrc_mlp* mlp);   // rebuild transposed weights after a parameter change
nrc_status launch_infer_fp32(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const uint32_t* countDev, size_t n,
                             float* rgb, cudaStream_t st);
// deferred != nullptr: the per-CTA partials are NOT reduced; deferred[0] = CTAs, deferred[1] = 1 if the histogram scatter was used --
// the caller finishes with launch_reduce_step (fused with the optimiser steps) or launch_reduce_partials
nrc_status launch_train_fp32(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const float* target, size_t n,
                             float lossScale, float* pred, cudaStream_t st, int* deferred = nullptr);
// per-launch scratch of both train kernels: partial arenas, the dL/dx0 rows, the per-CTA histograms (useHist: the shared-memory path applies)
nrc_status train_scratch(nrc_mlp* mlp, nrc_mrhe* m, size_t n, bool* useHist);
nrc_status launch_reduce_partials(nrc_mlp* mlp, nrc_mrhe* m, int blocks, bool hist, cudaStream_t st);
// steps.cu: {k_reduce_partials, k_nrc_step, k_mrhe_step} as one launch, where that is exactly what the three would do
bool reduce_step_fusable(const nrc_mlp* mlp, const nrc_mrhe* m, bool hist);
nrc_status launch_reduce_step(nrc_mlp* mlp, nrc_mrhe* m, int blocks, cudaStream_t st);
// steps.cu
nrc_status launch_nrc_step(nrc_mlp* mlp, cudaStream_t st);
nrc_status launch_mrhe_step(nrc_mrhe* m, cudaStream_t st);
nrc_status launch_init_normal(nrc_ctx* ctx, float* dst, size_t n, float scale, uint64_t seed);
// mlp_tc.cu (tcgen05 path)
nrc_status launch_train_tc(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const float* target, size_t n, float lossScale,
                           float* pred, cudaStream_t st, int* deferred = nullptr);
nrc_status launch_infer_tc(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const uint32_t* countDev, size_t n,
                           float* rgb, cudaStream_t st);
// split inference: k_encode_tiles (mlp_tc.cu) -> bf16 operand-image tiles -> k_mlp_stream (mlp_stream.cu)
nrc_status launch_encode_tiles(nrc_mrhe* m, const float* pos, const float* dir, const uint32_t* countDev, size_t n, void* featTiles,
                               cudaStream_t st);
nrc_status launch_mlp_stream(nrc_mlp* mlp, const void* featTiles, const uint32_t* countDev, size_t n, float* rgb, cudaStream_t st);
nrc_status launch_infer_split(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const uint32_t* countDev, size_t n, float* rgb,
                              cudaStream_t st);
