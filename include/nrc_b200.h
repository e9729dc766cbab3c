/*
 * nrc_b200.h -- C-ABI of the B200-native neural-radiance-cache hot path.
 *
 * This is the drop-in boundary for the hot path of MichalSzczekocki/NeuralRadianceCaching
 * (SURVEY.md section 8b): what the reference does through seven Vulkan descriptor sets and one
 * pre-recorded command buffer (src/NrcHpmRenderer.cpp:571-730), a host does here through plain
 * pointers and sizes.  No torch, no C++ types in the signatures.  Every entry point returns an
 * nrc_status; the text of the last failure on the calling thread is nrc_last_error().
 *
 * Pointer arguments marked [h|d] may be host or device memory; the library detects which
 * (cudaPointerGetAttributes) and copies host buffers to device scratch before it returns (pageable or page-locked alike: the
 * caller may reuse or free a host buffer as soon as the call returns).  All work is
 * issued on the context's stream; calls that return data to host memory synchronise that stream.
 *
 * There is NO CPU fallback: without a CUDA device nrc_ctx_create fails with NRC_ERR_NO_DEVICE.
 *
 * Parity preset: all quirks of the reference shaders are reproduced; the
 * transcendental functions that steer control flow are the ones in nrc_math.h.
 */
#ifndef NRC_B200_H
#define NRC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NRC_API __attribute__((visibility("default")))

typedef enum nrc_status {
    NRC_OK = 0,
    NRC_ERR_INVALID = 1,    /* bad argument */
    NRC_ERR_NO_DEVICE = 2,  /* no usable CUDA device / wrong architecture */
    NRC_ERR_CUDA = 3,       /* a CUDA runtime call failed; see nrc_last_error() */
    NRC_ERR_OOM = 4,
    NRC_ERR_OVERFLOW = 5    /* a device work pool overflowed (the call grows it and retries once) */
} nrc_status;

/* Packed MLP parameter arena, 20995 floats: W0..W4 (64x64, [row=out][col=in]), W5 (3x64),
 * b0..b4 (64 each), b5 (3).  Same order as set 3 bindings 0-5 / 18-23 of the reference
 * (data/shader/nrc-train/nrc-train.comp:40-128). */
#define NRC_MLP_PARAM_COUNT 20995
#define NRC_MLP_WEIGHT_COUNT 20672
#define NRC_MRHE_MAX_LEVELS 16
#define NRC_FEATURES 64 /* 32 MRHE + 32 OneBlob network inputs */

typedef struct nrc_ctx nrc_ctx;
typedef struct nrc_mlp nrc_mlp;           /* replaces en::NeuralRadianceCache (include/engine/graphics/NeuralRadianceCache.hpp:8-61) */
typedef struct nrc_mrhe nrc_mrhe;         /* replaces en::MRHE (include/engine/graphics/MRHE.hpp:7-47) */
typedef struct nrc_volume nrc_volume;     /* replaces vk::Texture3D + en::VolumeData (src/Texture3D.cpp:10-43, include/engine/objects/VolumeData.hpp:10-49) */
typedef struct nrc_envmap nrc_envmap;     /* replaces en::HdrEnvMap (include/engine/graphics/HdrEnvMap.hpp:8-72) */
typedef struct nrc_renderer nrc_renderer; /* replaces en::NrcHpmRenderer (include/engine/graphics/renderer/NrcHpmRenderer.hpp:17-96) */

/* ---- POD mirrors of the reference's uniform blocks ---- */

/* VolumeUniformData, include/engine/objects/VolumeData.hpp:10-19 */
typedef struct nrc_volume_params {
    float random[4];
    uint32_t useNN;
    uint32_t showNonNN;
    float densityFactor;
    float g;
    int32_t noNnSpp;
    int32_t withNnSpp;
} nrc_volume_params;

/* DirLightData, include/engine/graphics/DirLight.hpp:7-14 */
typedef struct nrc_dir_light {
    float color[3];
    float zenith;
    float dir[3];
    float azimuth;
    float strength;
} nrc_dir_light;

/* PointLight::UniformData, include/engine/graphics/PointLight.hpp:27-32 */
typedef struct nrc_point_light {
    float pos[3];
    float strength;
    float color[3];
} nrc_point_light;

/* CameraMatrices + camera position, include/engine/graphics/Camera.hpp:11-15 (column-major, glm) */
typedef struct nrc_camera {
    float projView[16];
    float invProjView[16];
    float pos[3];
} nrc_camera;

typedef enum nrc_frame_format {
    NRC_FRAME_RGBA32F = 0, /* vec4 outColor of nrc-forward.frag:837-870, 16 B/pixel */
    NRC_FRAME_BGRA8 = 1    /* VK_FORMAT_B8G8R8A8_UNORM colour attachment (src/VulkanAPI.cpp:325), 4 B/pixel */
} nrc_frame_format;

typedef enum nrc_resolution_mode {
    NRC_RES_INTENDED = 0,         /* tight-packed table of src/MRHE.cpp:111-121 */
    NRC_RES_AS_SHIPPED_STD140 = 1 /* what the shipped SPIR-V reads: res[1], res[5], res[9], res[13], then 0 */
} nrc_resolution_mode;

typedef enum nrc_mlp_precision {
    NRC_MLP_FP32 = 0,    /* CUDA-core fp32 kernels; the 1e-3 parity path */
    NRC_MLP_BF16_TC = 1  /* tcgen05 bf16 tensor-core kernels, fp32 accumulate; looser stated bound */
} nrc_mlp_precision;

/* How the bf16 tensor-core inference is organised (no reference counterpart: fwd.frag:503-534 is one function). */
typedef enum nrc_infer_mode {
    NRC_INFER_FUSED = 0, /* k_infer_tc: encode + six layers in one kernel per 128-query tile */
    NRC_INFER_SPLIT = 1  /* k_encode_tiles (gather kernel) -> bf16 operand-image tiles in HBM -> k_mlp_stream (pipelined tcgen05 MLP) */
} nrc_infer_mode;

/* ---- context ---- */
NRC_API const char* nrc_last_error(void);
NRC_API const char* nrc_version(void);
/* replaces VulkanAPI::Init + the per-class Init() statics (src/VulkanAPI.cpp:35-52) */
NRC_API nrc_status nrc_ctx_create(int device, nrc_ctx** out);
NRC_API nrc_status nrc_ctx_destroy(nrc_ctx* ctx);
NRC_API nrc_status nrc_sync(nrc_ctx* ctx);          /* vkQueueWaitIdle, src/main.cpp:217 */
NRC_API void* nrc_ctx_stream(nrc_ctx* ctx);         /* the cudaStream_t all work is issued on */
NRC_API int nrc_ctx_sm_count(nrc_ctx* ctx);
NRC_API nrc_status nrc_ctx_set_precision(nrc_ctx* ctx, nrc_mlp_precision p);
/* applies to NRC_MLP_BF16_TC inference (nrc_infer and the frame); default from the environment NRC_INFER_SPLIT=0/1 */
NRC_API nrc_status nrc_ctx_set_infer_mode(nrc_ctx* ctx, nrc_infer_mode mode);
/* number of kernels this context has launched so far (bench.py's gpu_launches) */
NRC_API uint64_t nrc_ctx_launch_count(nrc_ctx* ctx);

/* Per-kernel timing with CUDA events recorded on the launching stream (replaces nothing in the reference, which has
 * only a wall-clock FPS counter, src/Time.cpp:8-25).  While enabled every kernel launch of this context is
 * bracketed by two events; nrc_ctx_profile_read waits for them and returns accumulated milliseconds and launch
 * counts per kernel id (arrays of nrc_kernel_count() entries). */
/* Test hook, not a product path: exhaustive device-side comparison of the guard-free division / square-root sequences used on
 * the walks' serial chains (csrc/strict.cuh) with the IEEE routines they shortcut; mismatches[k] must all be 0. */
NRC_API nrc_status nrc_selftest_strict_math(nrc_ctx* ctx, uint64_t mismatches[4]);

NRC_API int nrc_kernel_count(void);
NRC_API const char* nrc_kernel_name(int kernelId);
NRC_API nrc_status nrc_ctx_profile_enable(nrc_ctx* ctx, int on);
NRC_API nrc_status nrc_ctx_profile_read(nrc_ctx* ctx, double* ms, uint64_t* counts, int reset);
/* the launches since the last read as (kernel id, start, end) in ms relative to the first launch: the timeline over both streams */
NRC_API nrc_status nrc_ctx_profile_timeline(nrc_ctx* ctx, int maxRecords, int* kernelId, float* startMs, float* endMs, int* count);

/* ---- NeuralRadianceCache: MLP parameter store (src/NeuralRadianceCache.cpp:370-512, 596-700) ---- */
/* seed != 0: W ~ N(0,1)*0.01 from a counter-based generator, biases/deltas/momenta zero. */
NRC_API nrc_status nrc_mlp_create(nrc_ctx* ctx, float learningRate, float weightDecay, float beta1, uint64_t seed, nrc_mlp** out);
NRC_API nrc_status nrc_mlp_destroy(nrc_mlp* mlp);
NRC_API nrc_status nrc_mlp_set_config(nrc_mlp* mlp, float learningRate, float weightDecay, float beta1);
NRC_API nrc_status nrc_mlp_set_params(nrc_mlp* mlp, const float* params /*[h|d] 20995*/);
NRC_API nrc_status nrc_mlp_set_state(nrc_mlp* mlp, const float* delta /*[h|d] or NULL*/, const float* momentum /*[h|d] or NULL*/);
NRC_API nrc_status nrc_mlp_get_params(nrc_mlp* mlp, float* params, float* delta, float* momentum /*[h|d], any may be NULL*/);
NRC_API nrc_status nrc_mlp_reset_stats(nrc_mlp* mlp);             /* ResetStats(), NeuralRadianceCache.cpp:541-545 */
NRC_API nrc_status nrc_mlp_get_stats(nrc_mlp* mlp, float* mseLoss /*host*/); /* GetStats(), :552-556 */
/* device address of the gradient arena handed to the collective: 20995 negative-gradient floats followed by
 * the mseLoss accumulator (20996 floats in all). */
/* GetStats() queued on the context's stream, no host wait: *mseLoss (page-locked host memory) is valid after a later
 * nrc_readback_wait / nrc_sync. */
NRC_API nrc_status nrc_mlp_get_stats_async(nrc_mlp* mlp, float* mseLoss);
NRC_API float* nrc_mlp_grad_ptr(nrc_mlp* mlp);
NRC_API size_t nrc_mlp_grad_count(nrc_mlp* mlp);

/* ---- MRHE: hash-table store (src/MRHE.cpp:79-225) ---- */
NRC_API nrc_status nrc_mrhe_resolutions(uint32_t minRes, uint32_t maxRes, uint32_t levelCount, nrc_resolution_mode mode, uint32_t out[16]);
/* resolutions == NULL -> the intended table.  seed != 0: tables ~ N(0,1)*0.1; delta tables zero. */
NRC_API nrc_status nrc_mrhe_create(nrc_ctx* ctx, float learningRate, float weightDecay, uint32_t levelCount, uint32_t hashTableSize,
                                   uint32_t featureCount, uint32_t minRes, uint32_t maxRes, const uint32_t* resolutions, uint64_t seed,
                                   nrc_mrhe** out);
NRC_API nrc_status nrc_mrhe_destroy(nrc_mrhe* mrhe);
NRC_API nrc_status nrc_mrhe_set_config(nrc_mrhe* mrhe, float learningRate, float weightDecay);
NRC_API size_t nrc_mrhe_table_bytes(nrc_mrhe* mrhe);               /* GetHashTableSize(), MRHE.cpp:267-270 (bytes) */
NRC_API nrc_status nrc_mrhe_set_tables(nrc_mrhe* mrhe, const float* tables /*[h|d] L*T*F*/, const float* delta /*or NULL*/);
NRC_API nrc_status nrc_mrhe_get_tables(nrc_mrhe* mrhe, float* tables, float* delta /*[h|d], may be NULL*/);
NRC_API float* nrc_mrhe_grad_ptr(nrc_mrhe* mrhe);                  /* device address of the delta tables */
/* floats at the head of the delta tables that the backward pass can touch: 2*T in the parity preset,
 * because the reference's scatter address has no level term (nrc-train.comp:977-978). */
NRC_API size_t nrc_mrhe_grad_extent(nrc_mrhe* mrhe);

/* ---- opt-in "fixed" preset (NOT what the reference ships -- every call below defaults to off) ----
 * nrc_mrhe_set_fixed_mode: EncodePosMrhe / BackpropMrhe (nrc-train.comp:266-339, 947-983) with the slips of SURVEY Appendix A2, A3, A9,
 *   A10 corrected: lerp factors are the grid-space fractions resPos - floor(resPos) and each blends its own axis; each of a level's two
 *   features gets its own error; gradients land in their level's table (nrc_mrhe_grad_extent becomes the whole table).
 * nrc_*_set_adam: nrc-step.comp / mrhe-step.comp with a second moment and bias correction (the shipped "Adam" is first-moment momentum
 *   only, A16): g' as shipped (scrub, weight decay); m = b1 m + (1-b1) g'; v = b2 v + (1-b2) g'^2; w += lr * mhat / (sqrt(vhat) + eps).
 *   The MLP keeps the beta1 of nrc_mlp_create; the tables get their own.  Switching on (or off) resets moments and step count.
 *   The fused data-parallel exchange (nrc_dp_*) applies the same Adam step inside its kernel (moments replicated, or sharded with the
 *   table). */
NRC_API nrc_status nrc_mrhe_set_fixed_mode(nrc_mrhe* mrhe, int on);
NRC_API nrc_status nrc_mlp_set_adam(nrc_mlp* mlp, int on, float beta2, float eps);
NRC_API nrc_status nrc_mrhe_set_adam(nrc_mrhe* mrhe, int on, float beta1, float beta2, float eps);
/* Adam moments and step counts, for checkpoints ([h|d] float arrays of 20 995 / L*T*F elements; any pointer may be NULL).  The MLP's
 * first moment is the reference's momentum arena (nrc_mlp_get_params / nrc_mlp_set_state).  NRC_ERR_INVALID unless Adam is on. */
NRC_API nrc_status nrc_mlp_get_adam_state(nrc_mlp* mlp, float* second, uint32_t* steps);
NRC_API nrc_status nrc_mlp_set_adam_state(nrc_mlp* mlp, const float* second, uint32_t steps);
NRC_API nrc_status nrc_mrhe_get_adam_state(nrc_mrhe* mrhe, float* m1, float* m2, uint32_t* steps);
NRC_API nrc_status nrc_mrhe_set_adam_state(nrc_mrhe* mrhe, const float* m1, const float* m2, uint32_t steps);

/* ---- stand-alone operators (what the reference runs inside its shaders; used by parity tests and bench) ---- */
/* EncodePosMrhe (nrc-train.comp:266-339): pos[N][3] -> feat[N][32]; idx[N][128] (level*8+corner) optional. */
NRC_API nrc_status nrc_encode(nrc_mrhe* mrhe, const float* pos, size_t n, float* feat, uint32_t* idx);
/* EncodeDirOneBlob (nrc-train.comp:352-365): dir[N][3] -> feat[N][32]. */
NRC_API nrc_status nrc_encode_dir(nrc_ctx* ctx, const float* dir, size_t n, float* feat);
/* Forward() (nrc-forward.frag:503-534): pos,dir[N][3] -> rgb[N][3]. */
NRC_API nrc_status nrc_infer(nrc_mlp* mlp, nrc_mrhe* mrhe, const float* pos, const float* dir, size_t n, float* rgb);
/* The two halves of Forward() as separate operators (NRC_INFER_SPLIT): EncodeRay (nrc-train.comp:703-713) into bf16
 * operand-image tiles -- 16 KB per 128 queries, 16-byte chunk (kc, row) at (kc*128 + row)*16, kc = 8 consecutive network
 * inputs, rows past n zero -- and ApplyWeights0..5 / ActivateNr1..6 (nrc-train.comp:547-701) from such tiles.
 * featTiles is DEVICE memory of nrc_feature_tile_bytes(n) bytes; pos/dir/rgb are [h|d]. */
NRC_API size_t nrc_feature_tile_bytes(size_t n);
NRC_API nrc_status nrc_encode_tiles(nrc_mrhe* mrhe, const float* pos, const float* dir, size_t n, void* featTiles);
NRC_API nrc_status nrc_mlp_forward_tiles(nrc_mlp* mlp, const void* featTiles, size_t n, float* rgb);
/* Backprop() (nrc-train.comp:985-1028) over N records: adds the negative gradients (times lossScale =
 * ONE_OVER_PIXEL_COUNT, nrc-train.comp:210) into the delta arenas and the mean-square error into mseLoss.
 * pred[N][3] (optional) receives the network outputs. */
NRC_API nrc_status nrc_train_accumulate(nrc_mlp* mlp, nrc_mrhe* mrhe, const float* pos, const float* dir, const float* target, size_t n,
                                        float lossScale, float* pred);
/* No counterpart in the reference.  nrc_train_accumulate for callers that cannot call
 * cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it): always the fp32 kernel, pos / dir / target / pred
 * must be DEVICE memory, n <= nrc_train_accumulate_static_max_samples() (8192; larger n is rejected with NRC_ERR_INVALID).  Calls no
 * cudaMalloc / cudaFree: its scratch is one file-scope __device__ array (~3.4 MB per GPU, held until the process exits) shared by every
 * call -- calls whose kernels overlap on one GPU (different contexts / streams) race on it; calls on one context are stream-ordered. */
NRC_API size_t nrc_train_accumulate_static_max_samples(void);
NRC_API nrc_status nrc_train_accumulate_static(nrc_mlp* mlp, nrc_mrhe* mrhe, const float* pos, const float* dir, const float* target, size_t n,
                                               float lossScale, float* pred);
/* nrc-step.comp:394-406 then mrhe-step.comp:63-74 (either handle may be NULL to skip it). */
NRC_API nrc_status nrc_optimizer_step(nrc_mlp* mlp, nrc_mrhe* mrhe);

/* ---- scene objects ---- */
/* density: raw float32 grid in the file order of ReadFileDensity3D (src/read_file.cpp:61-74),
 * index = x*ny*nz + y*nz + z; quantised to UNORM8 like src/Texture3D.cpp:25-40. */
NRC_API nrc_status nrc_volume_create(nrc_ctx* ctx, const float* density /*host*/, uint32_t nx, uint32_t ny, uint32_t nz, nrc_volume** out);
NRC_API nrc_status nrc_volume_destroy(nrc_volume* vol);
NRC_API nrc_status nrc_volume_set_params(nrc_volume* vol, const nrc_volume_params* p);   /* VolumeData::Update, src/VolumeData.cpp:105-109 */
/* getDensity() (nrc-forward.frag:574-582) at N world positions, for parity tests. */
NRC_API nrc_status nrc_volume_sample(nrc_volume* vol, const float* pos, size_t n, float* density);

/* rgba: float RGBA, row 0 first (already flipped as stbi does for the reference, src/read_file.cpp:83);
 * invCdfX[h*w], invCdfY[h] as Hdr4fToCdf returns them (may be NULL: unused by the live code path). */
NRC_API nrc_status nrc_envmap_create(nrc_ctx* ctx, uint32_t width, uint32_t height, const float* rgba, const float* invCdfX,
                                     const float* invCdfY, nrc_envmap** out);
NRC_API nrc_status nrc_envmap_destroy(nrc_envmap* env);
NRC_API nrc_status nrc_envmap_set_strengths(nrc_envmap* env, float directStrength, float hpmStrength);  /* HdrEnvMap.cpp:108, 259-279 */
/* Opt-in estimator, not the shipped behaviour: `halfSampleCount = sampleCount / 2` as left commented out at
 * nrc-forward.frag:715 -- 4 of the 8 env directions per lit vertex are drawn from the map's inverse CDFs
 * (nrc-forward.frag:727-744).  Needs an env map created with both inverse-CDF images. */
NRC_API nrc_status nrc_envmap_set_importance_sampling(nrc_envmap* env, int on);
/* Hdr4fToCdf (src/read_file.cpp:141-206) on the device. */
NRC_API nrc_status nrc_hdr4f_to_cdf(nrc_ctx* ctx, const float* rgba, uint32_t width, uint32_t height, float* invCdfX, float* invCdfY);

/* ---- NrcHpmRenderer (src/NrcHpmRenderer.cpp) ---- */
NRC_API nrc_status nrc_renderer_create(nrc_ctx* ctx, uint32_t width, uint32_t height, uint32_t trainWidth, uint32_t trainHeight,
                                       nrc_volume* vol, nrc_envmap* env, nrc_mlp* mlp, nrc_mrhe* mrhe, nrc_renderer** out);
NRC_API nrc_status nrc_renderer_destroy(nrc_renderer* r);
NRC_API nrc_status nrc_renderer_set_camera(nrc_renderer* r, const nrc_camera* cam);        /* Camera::UpdateUniformBuffer, Camera.cpp:164-173 */
NRC_API nrc_status nrc_renderer_set_dir_light(nrc_renderer* r, const nrc_dir_light* l);
NRC_API nrc_status nrc_renderer_set_point_light(nrc_renderer* r, const nrc_point_light* l);
/* Multi-GPU partition: this rank renders frame rows [y0,y1) and traces train invocations [i0,i1)
 * (i = y*trainWidth + x).  Seeds use GLOBAL coordinates, so the union over ranks equals one GPU. */
NRC_API nrc_status nrc_renderer_set_partition(nrc_renderer* r, uint32_t y0, uint32_t y1, uint32_t i0, uint32_t i1);
/* Interleaved partition for load balance (cloud pixels cluster mid-frame): this rank renders rows yFirst, yFirst+yStride, ...
 * and traces train invocations iFirst, iFirst+iStride, ...  (rank, worldSize, rank, worldSize) for data-parallel hosts. */
NRC_API nrc_status nrc_renderer_set_partition_strided(nrc_renderer* r, uint32_t yFirst, uint32_t yStride, uint32_t iFirst, uint32_t iStride);
/* Capacity of the render vertex pool while the cache is on.  The reference's bound is 32 lit vertices per path
 * (nrc-forward.frag:772-806); a pool of that size cannot overflow.  0 = the library chooses: the worst-case pool when
 * its three sets are at most a sixth of the GPU's memory (25.5 GB for a 1920x1080 frame at 1 spp), else 4 vertices per
 * path.  A frame that overflows a smaller pool is traced again with the worst-case pool before any read-back or
 * nrc_sync returns; the price of a smaller pool is one host wait per frame for the overflow flag. */
NRC_API nrc_status nrc_renderer_set_pool_capacity(nrc_renderer* r, uint64_t vertices);
/* the capacity the last frame was traced with (after the library's choice or a growth) */
NRC_API nrc_status nrc_renderer_get_pool_capacity(nrc_renderer* r, uint64_t* vertices);
/* Opt-in estimator, not the shipped behaviour: the env-map term of TraceScene in the training targets, commented out at
 * nrc-train.comp:1118 (SampleHdrEnvMap(pos, dir, 8) as in nrc-forward.frag:751-755).  Changes the train walk's RNG
 * consumption (16 more draws per lit vertex, before the two of NewRayDir). */
NRC_API nrc_status nrc_renderer_set_train_env_light(nrc_renderer* r, int on);
/* ResizeFrame, NrcHpmRenderer.cpp:103-126 */
NRC_API nrc_status nrc_resize(nrc_renderer* r, uint32_t width, uint32_t height);

/* Render(): the whole recorded command buffer -- train, nrc-step, mrhe-step, render (asynchronous). */
NRC_API nrc_status nrc_render(nrc_renderer* r);
/* The same frame in the three pieces a data-parallel host needs (gradient all-reduce goes between 1 and 2):
 *   1. nrc_frame_train     : nrc-train.comp over this rank's invocations (accumulates gradients)
 *   2. nrc_optimizer_step  : nrc-step.comp + mrhe-step.comp
 *   3. nrc_frame_render    : nrc-forward.{vert,frag} over this rank's rows */
NRC_API nrc_status nrc_frame_train(nrc_renderer* r);
NRC_API nrc_status nrc_frame_render(nrc_renderer* r);

NRC_API size_t nrc_frame_bytes(nrc_renderer* r, nrc_frame_format fmt);   /* GetImageDataSize(): w*h*4 for BGRA8 */
NRC_API void* nrc_frame_ptr(nrc_renderer* r, nrc_frame_format fmt);      /* device address of the full frame */
NRC_API nrc_status nrc_read_frame(nrc_renderer* r, void* dst /*[h|d]*/, nrc_frame_format fmt);

/* this rank's rows only, compacted (row k of dst = frame row y0 + k*stride); nrc_frame_partition_bytes is its size */
NRC_API nrc_status nrc_read_frame_partition(nrc_renderer* r, void* dst /*[h|d]*/, nrc_frame_format fmt);
NRC_API size_t nrc_frame_partition_bytes(nrc_renderer* r, nrc_frame_format fmt);
/* The same copy without the host wait (the reference presents frame k while it records frame k+1, src/main.cpp:197-230):
 * queued behind the frame just rendered, on a copy stream; dst (page-locked host or device memory) is valid after
 * nrc_readback_wait(ticket).  Up to 4 read-backs may be outstanding. */
NRC_API nrc_status nrc_read_frame_partition_async(nrc_renderer* r, void* dst, nrc_frame_format fmt, int* ticket);
NRC_API nrc_status nrc_readback_wait(nrc_renderer* r, int ticket);

/* ---- data-parallel exchange fused with the optimiser steps (the reference is single-GPU) ----
 * One process per GPU.  Every rank creates an nrc_dp for its replica, publishes nrc_dp_handle_bytes() opaque bytes
 * (nrc_dp_export; CUDA-IPC handles of its gradient arenas and flag block), gathers the handles of all ranks in rank
 * order by whatever transport the host has (torch.distributed, MPI, a file) and calls nrc_dp_connect.  After that
 * nrc_dp_reduce_step replaces {all-reduce(MLP gradients), all-reduce(delta-table head), nrc-step, mrhe-step} of
 * one frame by one kernel per rank that reads the peers' gradients over NVLink, sums them in rank order and
 * applies both steps (src/NeuralRadianceCache.cpp:512-539 and src/MRHE.cpp:189-225 are the single-GPU steps it
 * subsumes).  Every rank must call it once per frame; the kernels wait for each other on the device. */
typedef struct nrc_dp nrc_dp;
NRC_API size_t nrc_dp_handle_bytes(void);
NRC_API nrc_status nrc_dp_create(nrc_mlp* mlp, nrc_mrhe* mrhe, int rank, int world, nrc_dp** out);
NRC_API nrc_status nrc_dp_export(nrc_dp* dp, void* handle /*host, nrc_dp_handle_bytes()*/);
NRC_API nrc_status nrc_dp_connect(nrc_dp* dp, const void* allHandles /*host, world handles in rank order*/, int sameProcess);
NRC_API nrc_status nrc_dp_reduce_step(nrc_dp* dp);
/* replicas inside one process (single-GPU parity test): phase 0 on every replica, then phase 1 on every replica.  Phase 2 is a test
 * hook: this rank posts its gradients and then waits for peers that never arrive, to exercise the time-out (fails closed). */
NRC_API nrc_status nrc_dp_reduce_step_phase(nrc_dp* dp, int phase);
/* 1 if a handshake gave up waiting for a peer (about 30 s; a crashed rank must not hang the other GPUs).  The exchange FAILS CLOSED:
 * the block that gave up skips its step and its clear, the mark is written into every rank's flag block, and from then on
 * nrc_dp_reduce_step, nrc_sync and nrc_readback_wait of every rank return NRC_ERR_CUDA. */
NRC_API nrc_status nrc_dp_status(nrc_dp* dp, int* timedOut);
NRC_API nrc_status nrc_dp_destroy(nrc_dp* dp);

/* Parity probes of the last frame (rows/invocations of this rank's partition only are meaningful):
 *   counters[h*w][4] = {randDraws, litVertices, walkSteps, queryEmitted}
 *   query[h*w][8]    = {pos.xyz, dir.xyz, weight, valid}
 *   records[n][9]    = {pos.xyz, dir.xyz, target.rgb}, trainCounters[n][4] as above (n = trainW*trainH) */
NRC_API nrc_status nrc_read_render_probes(nrc_renderer* r, uint32_t* counters, float* query);
NRC_API nrc_status nrc_read_train_probes(nrc_renderer* r, float* records, uint32_t* trainCounters);
/* queries inferred + samples trained by the last frame on this rank, and density taps taken */
NRC_API nrc_status nrc_frame_stats(nrc_renderer* r, uint64_t* inferQueries, uint64_t* trainSamples, uint64_t* densityTaps);
/* running totals over all frames since creation (or the last reset) on this rank:
 * totals[0] inference queries, [1] training samples, [2] render-side density taps, [3] train-side density taps,
 * [4] render vertices lit, [5] train vertices lit.  Synchronises the context. */
NRC_API nrc_status nrc_frame_totals(nrc_renderer* r, uint64_t totals[6], int reset);

#ifdef __cplusplus
}
#endif
#endif /* NRC_B200_H */
