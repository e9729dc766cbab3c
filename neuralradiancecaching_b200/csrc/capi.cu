// capi.cu -- extern "C" entry points of include/nrc_b200.h (context, parameter stores, stand-alone operators,
// scene objects).  The renderer's entry points live in renderer.cu.
#include <stdarg.h>
#include <string.h>
#include <math.h>
#include "nrc_internal.cuh"

// ---- errors -------------------------------------------------------------------------------------------------
static thread_local char g_err[1024] = "";

void nrc_set_error(const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* nrc_last_error(void) { return g_err; }
extern "C" const char* nrc_version(void) { return "nrc_b200 0.1 (sm_100a)"; }

// ---- staging helpers ----------------------------------------------------------------------------------------
bool nrc_is_device_ptr(const void* p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

nrc_status nrc_scratch(nrc_ctx* ctx, int slot, size_t bytes, void** out)
{
    if (ctx->scratchBytes[slot] < bytes) {
        NRC_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ctx->scratch[slot]) NRC_CUDA(cudaFree(ctx->scratch[slot]));
        ctx->scratch[slot] = nullptr;
        ctx->scratchBytes[slot] = 0;
        size_t cap = bytes + bytes / 4 + 256;
        NRC_CUDA(cudaMalloc(&ctx->scratch[slot], cap));
        ctx->scratchBytes[slot] = cap;
    }
    *out = ctx->scratch[slot];
    return NRC_OK;
}

nrc_status nrc_stage_in(nrc_ctx* ctx, int slot, const void* p, size_t bytes, const void** dev)
{
    if (bytes == 0 || nrc_is_device_ptr(p)) {
        *dev = p;
        return NRC_OK;
    }
    void* d;
    NRC_TRY(nrc_scratch(ctx, slot, bytes, &d));
    // pageable host memory is staged by the runtime before cudaMemcpyAsync returns; a PAGE-LOCKED source is read by the DMA engine
    // later, so wait for it: the caller may reuse or free the buffer as soon as the entry point returns
    NRC_CUDA(cudaMemcpyAsync(d, p, bytes, cudaMemcpyHostToDevice, ctx->stream));
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) cudaGetLastError();
    else if (at.type == cudaMemoryTypeHost) NRC_CUDA(cudaStreamSynchronize(ctx->stream));
    *dev = d;
    return NRC_OK;
}

nrc_status nrc_stage_out(nrc_ctx* ctx, int slot, void* dst, size_t bytes, void** dev)
{
    if (bytes == 0 || nrc_is_device_ptr(dst)) {
        *dev = dst;
        return NRC_OK;
    }
    return nrc_scratch(ctx, slot, bytes, dev);
}

nrc_status nrc_stage_out_finish(nrc_ctx* ctx, void* dst, const void* dev, size_t bytes)
{
    if (bytes == 0 || dst == dev) return NRC_OK;
    NRC_CUDA(cudaMemcpyAsync(dst, dev, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream));
    return NRC_OK;
}

static nrc_status copy_in(nrc_ctx* ctx, void* devDst, const void* src, size_t bytes)
{
    cudaMemcpyKind k = nrc_is_device_ptr(src) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    NRC_CUDA(cudaMemcpyAsync(devDst, src, bytes, k, ctx->stream));
    if (k == cudaMemcpyHostToDevice) NRC_CUDA(cudaStreamSynchronize(ctx->stream));   // caller may free src on return
    return NRC_OK;
}

static nrc_status copy_out(nrc_ctx* ctx, void* dst, const void* devSrc, size_t bytes)
{
    cudaMemcpyKind k = nrc_is_device_ptr(dst) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    NRC_CUDA(cudaMemcpyAsync(dst, devSrc, bytes, k, ctx->stream));
    if (k == cudaMemcpyDeviceToHost) NRC_CUDA(cudaStreamSynchronize(ctx->stream));
    return NRC_OK;
}

// ---- per-kernel event timing ------------------------------------------------------------------------------
static const char* const g_kernelNames[K_COUNT] = {
    "k_encode_pos", "k_encode_dir", "k_infer_fp32", "k_train_fp32", "k_reduce_partials", "k_transpose_weights", "k_nrc_step",
    "k_mrhe_step", "k_init_normal", "k_render_walk", "k_train_walk", "k_light<render>", "k_light<train>", "k_render_combine",
    "k_train_combine", "k_composite", "k_build_bricks", "k_volume_sample", "k_pack_records", "k_cdf_rows", "k_cdf_y",
    "k_infer_tc", "k_train_tc", "k_pack_bf16", "misc", "k_dp_reduce_step", "k_encode_tiles", "k_mlp_stream", "k_render_walk_tail", "k_reduce_step"};

void nrc_prof_begin(nrc_ctx* ctx, cudaStream_t st, int kid)
{
    if (!ctx->profiling) return;
    NrcProfRec r;
    if (!ctx->profFree.empty()) {
        r = ctx->profFree.back();
        ctx->profFree.pop_back();
    } else {
        if (cudaEventCreate(&r.a) != cudaSuccess || cudaEventCreate(&r.b) != cudaSuccess) { cudaGetLastError(); return; }
    }
    r.kid = kid;
    cudaEventRecord(r.a, st);
    ctx->profOpen.push_back(r);
}

void nrc_prof_end(nrc_ctx* ctx, cudaStream_t st)
{
    if (!ctx->profiling || ctx->profOpen.empty()) return;
    cudaEventRecord(ctx->profOpen.back().b, st);
}

static void prof_drain(nrc_ctx* c)
{
    for (NrcProfRec& r : c->profOpen) {
        float ms = 0.0f;
        if (cudaEventSynchronize(r.b) == cudaSuccess && cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
            c->profMs[r.kid] += ms;
            c->profCount[r.kid]++;
        } else {
            cudaGetLastError();
        }
        c->profFree.push_back(r);
    }
    c->profOpen.clear();
}

// The launches recorded since the last read, in launch order, as (kernel id, start ms, end ms) relative to the first
// one's start: the frame's timeline across both streams.  Consumes the records (they are also accumulated).
extern "C" nrc_status nrc_ctx_profile_timeline(nrc_ctx* c, int maxRecords, int* kid, float* startMs, float* endMs, int* count)
{
    NRC_REQUIRE(c && kid && startMs && endMs && count, "null argument");
    int n = 0;
    if (!c->profOpen.empty()) {
        cudaEvent_t ref = c->profOpen.front().a;
        for (NrcProfRec& r : c->profOpen) {
            if (n >= maxRecords) break;
            float s = 0.0f, e = 0.0f;
            if (cudaEventSynchronize(r.b) != cudaSuccess || cudaEventElapsedTime(&s, ref, r.a) != cudaSuccess ||
                cudaEventElapsedTime(&e, ref, r.b) != cudaSuccess) { cudaGetLastError(); continue; }
            kid[n] = r.kid; startMs[n] = s; endMs[n] = e;
            n++;
        }
    }
    *count = n;
    prof_drain(c);
    return NRC_OK;
}

extern "C" int nrc_kernel_count(void) { return K_COUNT; }
extern "C" const char* nrc_kernel_name(int kid) { return (kid >= 0 && kid < K_COUNT) ? g_kernelNames[kid] : ""; }

extern "C" nrc_status nrc_ctx_profile_enable(nrc_ctx* c, int on)
{
    NRC_REQUIRE(c, "null context");
    prof_drain(c);
    c->profiling = (on != 0);
    return NRC_OK;
}

extern "C" nrc_status nrc_ctx_profile_read(nrc_ctx* c, double* ms, uint64_t* counts, int reset)
{
    NRC_REQUIRE(c, "null context");
    prof_drain(c);
    for (int k = 0; k < K_COUNT; k++) {
        if (ms) ms[k] = c->profMs[k];
        if (counts) counts[k] = c->profCount[k];
        if (reset) { c->profMs[k] = 0.0; c->profCount[k] = 0; }
    }
    return NRC_OK;
}

// ---- context -------------------------------------------------------------------------------------------------
extern "C" nrc_status nrc_ctx_create(int device, nrc_ctx** out)
{
    if (!out) return NRC_ERR_INVALID;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        nrc_set_error("no CUDA device: this library has no CPU fallback");
        return NRC_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= count) {
        nrc_set_error("device %d out of range (have %d)", device, count);
        return NRC_ERR_INVALID;
    }
    cudaDeviceProp prop;
    NRC_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        nrc_set_error("device %d is sm_%d%d; this library is built for sm_100a (B200) only", device, prop.major, prop.minor);
        return NRC_ERR_NO_DEVICE;
    }
    NRC_CUDA(cudaSetDevice(device));
    nrc_ctx* c = new nrc_ctx();
    c->device = device;
    c->smCount = prop.multiProcessorCount;
    // the main stream carries the frame's critical path (the latency-bound train chain): highest priority, so its blocks
    // are placed before pending blocks of the render walk running beside it on stream2
    int prLow = 0, prHigh = 0;
    NRC_CUDA(cudaDeviceGetStreamPriorityRange(&prLow, &prHigh));
    NRC_CUDA(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prHigh));
    NRC_CUDA(cudaStreamCreateWithPriority(&c->stream2, cudaStreamNonBlocking, prLow));
    // the train trace is the long pole of the frame: just below the main stream (whose kernels are short), above the render trace
    NRC_CUDA(cudaStreamCreateWithPriority(&c->stream3, cudaStreamNonBlocking, (prHigh + 1 <= prLow) ? prHigh + 1 : prHigh));
    c->streamWalk[0] = c->stream3;
    for (int i = 1; i < NRC_TRAIN_WALK_STREAMS; i++)
        NRC_CUDA(cudaStreamCreateWithPriority(&c->streamWalk[i], cudaStreamNonBlocking, (prHigh + 1 <= prLow) ? prHigh + 1 : prHigh));
    NRC_CUDA(cudaEventCreateWithFlags(&c->evTrainTrace, cudaEventDisableTiming));
    NRC_CUDA(cudaStreamCreateWithPriority(&c->stream4, cudaStreamNonBlocking, (prHigh + 1 <= prLow) ? prHigh + 1 : prHigh));
    NRC_CUDA(cudaStreamCreateWithFlags(&c->streamCopy, cudaStreamNonBlocking));
    NRC_CUDA(cudaStreamCreateWithPriority(&c->stream5, cudaStreamNonBlocking, prLow));
    NRC_CUDA(cudaEventCreateWithFlags(&c->evRenderWalk, cudaEventDisableTiming));
    NRC_CUDA(cudaEventCreateWithFlags(&c->evTrainWalk, cudaEventDisableTiming));
    for (int i = 0; i < NRC_TRAIN_SETS; i++) NRC_CUDA(cudaEventCreateWithFlags(&c->evTrainFree[i], cudaEventDisableTiming));
    NRC_CUDA(cudaEventCreateWithFlags(&c->evJoin, cudaEventDisableTiming));
    if (const char* e = getenv("NRC_INFER_SPLIT")) c->inferMode = atoi(e) ? NRC_INFER_SPLIT : NRC_INFER_FUSED;
    *out = c;
    return NRC_OK;
}

extern "C" nrc_status nrc_ctx_destroy(nrc_ctx* c)
{
    if (!c) return NRC_OK;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    cudaStreamSynchronize(c->stream2);
    for (int i = 0; i < NRC_TRAIN_WALK_STREAMS; i++) cudaStreamSynchronize(c->streamWalk[i]);
    cudaStreamSynchronize(c->stream4);
    cudaStreamSynchronize(c->stream5);
    cudaStreamSynchronize(c->streamCopy);
    for (int i = 0; i < 6; i++) if (c->scratch[i]) cudaFree(c->scratch[i]);
    if (c->pinned) cudaFreeHost(c->pinned);
    prof_drain(c);
    for (NrcProfRec& r : c->profFree) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
    cudaEventDestroy(c->evJoin);
    cudaStreamDestroy(c->stream);
    cudaStreamDestroy(c->stream2);
    for (int i = 0; i < NRC_TRAIN_WALK_STREAMS; i++) cudaStreamDestroy(c->streamWalk[i]);
    cudaEventDestroy(c->evTrainTrace);
    cudaEventDestroy(c->evTrainWalk);
    for (int i = 0; i < NRC_TRAIN_SETS; i++) cudaEventDestroy(c->evTrainFree[i]);
    cudaStreamDestroy(c->stream4);
    cudaStreamDestroy(c->stream5);
    cudaEventDestroy(c->evRenderWalk);
    cudaStreamDestroy(c->streamCopy);
    delete c;
    return NRC_OK;
}

nrc_status nrc_ctx_check_dp(nrc_ctx* c)
{
    for (size_t i = 0; i < c->dpFailWords.size(); i++)
        if (*c->dpFailWords[i] != 0u) {
            nrc_set_error("a fused data-parallel exchange timed out waiting for a peer rank and failed closed (step and clear skipped); "
                          "the replicas have diverged -- restart the job");
            return NRC_ERR_CUDA;
        }
    return NRC_OK;
}

extern "C" nrc_status nrc_sync(nrc_ctx* c)
{
    NRC_REQUIRE(c, "null context");
    NRC_CUDA(cudaStreamSynchronize(c->stream));
    NRC_TRY(nrc_ctx_check_dp(c));
    // a frame traced with a vertex pool below the worst case is checked now (and redone if it overflowed)
    for (size_t i = 0; i < c->renderers.size(); i++) NRC_TRY(nrc_renderer_resolve_pending(c->renderers[i]));
    NRC_CUDA(cudaStreamSynchronize(c->stream));
    return NRC_OK;
}

extern "C" void* nrc_ctx_stream(nrc_ctx* c) { return c ? (void*)c->stream : nullptr; }
extern "C" int nrc_ctx_sm_count(nrc_ctx* c) { return c ? c->smCount : 0; }
extern "C" uint64_t nrc_ctx_launch_count(nrc_ctx* c) { return c ? c->launches : 0; }

extern "C" nrc_status nrc_ctx_set_precision(nrc_ctx* c, nrc_mlp_precision p)
{
    NRC_REQUIRE(c, "null context");
    NRC_REQUIRE(p == NRC_MLP_FP32 || p == NRC_MLP_BF16_TC, "unknown precision");
    c->precision = p;
    return NRC_OK;
}

extern "C" nrc_status nrc_ctx_set_infer_mode(nrc_ctx* c, nrc_infer_mode mode)
{
    NRC_REQUIRE(c, "null context");
    NRC_REQUIRE(mode == NRC_INFER_FUSED || mode == NRC_INFER_SPLIT, "unknown inference mode");
    c->inferMode = mode;
    return NRC_OK;
}

// ---- MLP parameter store --------------------------------------------------------------------------------------
extern "C" nrc_status nrc_mlp_create(nrc_ctx* ctx, float lr, float wd, float beta1, uint64_t seed, nrc_mlp** out)
{
    NRC_REQUIRE(ctx && out, "null argument");
    NRC_CUDA(cudaSetDevice(ctx->device));
    nrc_mlp* m = new nrc_mlp();
    m->ctx = ctx; m->lr = lr; m->wd = wd; m->beta1 = beta1;
    const size_t padded = 21504 * sizeof(float);
    NRC_CUDA(cudaMalloc(&m->params, padded));
    NRC_CUDA(cudaMalloc(&m->grad, padded));
    NRC_CUDA(cudaMalloc(&m->momentum, padded));
    NRC_CUDA(cudaMalloc(&m->paramsT, padded));
    NRC_CUDA(cudaMemsetAsync(m->params, 0, padded, ctx->stream));
    NRC_CUDA(cudaMemsetAsync(m->grad, 0, padded, ctx->stream));
    NRC_CUDA(cudaMemsetAsync(m->momentum, 0, padded, ctx->stream));
    NRC_CUDA(cudaMemsetAsync(m->paramsT, 0, padded, ctx->stream));
    if (seed != 0) NRC_TRY(launch_init_normal(ctx, m->params, NRC_MLP_WEIGHT_COUNT, 0.01f, seed));   // NeuralRadianceCache.cpp:640-646
    m->dirtyT = true;
    *out = m;
    return NRC_OK;
}

extern "C" nrc_status nrc_mlp_destroy(nrc_mlp* m)
{
    if (!m) return NRC_OK;
    cudaStreamSynchronize(m->ctx->stream);
    cudaFree(m->params); cudaFree(m->grad); cudaFree(m->momentum); cudaFree(m->paramsT);
    if (m->partial) cudaFree(m->partial);
    if (m->gScratch) cudaFree(m->gScratch);
    if (m->histPartial) cudaFree(m->histPartial);
    if (m->paramsBf16) cudaFree(m->paramsBf16);
    if (m->featTiles) cudaFree(m->featTiles);
    if (m->second) cudaFree(m->second);
    delete m;
    return NRC_OK;
}

extern "C" nrc_status nrc_mlp_set_config(nrc_mlp* m, float lr, float wd, float beta1)
{
    NRC_REQUIRE(m, "null mlp");
    m->lr = lr; m->wd = wd; m->beta1 = beta1;
    return NRC_OK;
}

extern "C" nrc_status nrc_mlp_set_params(nrc_mlp* m, const float* params)
{
    NRC_REQUIRE(m && params, "null argument");
    NRC_TRY(copy_in(m->ctx, m->params, params, NRC_MLP_PARAM_COUNT * sizeof(float)));
    m->dirtyT = true;
    m->dirtyBf16 = true;
    return NRC_OK;
}

extern "C" nrc_status nrc_mlp_set_state(nrc_mlp* m, const float* delta, const float* momentum)
{
    NRC_REQUIRE(m, "null mlp");
    if (delta) NRC_TRY(copy_in(m->ctx, m->grad, delta, NRC_MLP_PARAM_COUNT * sizeof(float)));
    if (momentum) NRC_TRY(copy_in(m->ctx, m->momentum, momentum, NRC_MLP_PARAM_COUNT * sizeof(float)));
    return NRC_OK;
}

extern "C" nrc_status nrc_mlp_get_params(nrc_mlp* m, float* params, float* delta, float* momentum)
{
    NRC_REQUIRE(m, "null mlp");
    if (params) NRC_TRY(copy_out(m->ctx, params, m->params, NRC_MLP_PARAM_COUNT * sizeof(float)));
    if (delta) NRC_TRY(copy_out(m->ctx, delta, m->grad, NRC_MLP_PARAM_COUNT * sizeof(float)));
    if (momentum) NRC_TRY(copy_out(m->ctx, momentum, m->momentum, NRC_MLP_PARAM_COUNT * sizeof(float)));
    return NRC_OK;
}

extern "C" nrc_status nrc_mlp_reset_stats(nrc_mlp* m)
{
    NRC_REQUIRE(m, "null mlp");
    NRC_CUDA(cudaMemsetAsync(m->grad + NRC_MLP_PARAM_COUNT, 0, 2 * sizeof(float), m->ctx->stream));   // + the exchange's running total
    return NRC_OK;
}

extern "C" nrc_status nrc_mlp_get_stats(nrc_mlp* m, float* mseLoss)
{
    NRC_REQUIRE(m && mseLoss, "null argument");
    float v[2] = {0.0f, 0.0f};   // [0] accumulated on this GPU since the last exchange, [1] global total published by the fused exchange (0 without one)
    NRC_CUDA(cudaMemcpyAsync(v, m->grad + NRC_MLP_PARAM_COUNT, 2 * sizeof(float), cudaMemcpyDeviceToHost, m->ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(m->ctx->stream));
    *mseLoss = v[0] + v[1];
    return NRC_OK;
}

// GetStats() without the host wait: the copy is queued on the main stream; dst (page-locked host memory) holds the value
// once a later nrc_readback_wait / nrc_sync has returned.
extern "C" nrc_status nrc_mlp_get_stats_async(nrc_mlp* m, float* mseLoss)
{
    NRC_REQUIRE(m && mseLoss, "null argument");
    // with a fused exchange attached the frame's loss sits in the running-total slot (the exchanged slot has been cleared by then)
    NRC_CUDA(cudaMemcpyAsync(mseLoss, m->grad + NRC_MLP_PARAM_COUNT + (m->lossTotalSlot ? 1 : 0), sizeof(float), cudaMemcpyDeviceToHost,
                             m->ctx->stream));
    return NRC_OK;
}

extern "C" float* nrc_mlp_grad_ptr(nrc_mlp* m) { return m ? m->grad : nullptr; }
extern "C" size_t nrc_mlp_grad_count(nrc_mlp* m) { return m ? (size_t)NRC_MLP_PARAM_COUNT + 1 : 0; }

// ---- MRHE -----------------------------------------------------------------------------------------------------
extern "C" nrc_status nrc_mrhe_resolutions(uint32_t minRes, uint32_t maxRes, uint32_t levelCount, nrc_resolution_mode mode, uint32_t out[16])
{
    NRC_REQUIRE(out && levelCount >= 2 && levelCount <= 16 && minRes > 0 && maxRes >= minRes, "bad resolution arguments");
    // src/MRHE.cpp:111-121, float32 arithmetic with the C library's expf/logf/powf like the reference
    uint32_t tight[16];
    float b = expf((logf((float)maxRes) - logf((float)minRes)) / (float)(levelCount - 1));
    for (uint32_t i = 0; i < 16; i++) tight[i] = 0;
    for (uint32_t i = 0; i < levelCount; i++) tight[i] = (uint32_t)((float)minRes * powf(b, (float)i));
    if (mode == NRC_RES_INTENDED) {
        for (int i = 0; i < 16; i++) out[i] = tight[i];
    } else if (mode == NRC_RES_AS_SHIPPED_STD140) {
        // the shader's std140 block has resolutions at byte 32, stride 16; the host packs them at byte 28, stride 4
        // => shader res[i] = host word (32 + 16*i - 28)/4 = 1 + 4*i, and reads past the 92-byte block give 0 here
        for (int i = 0; i < 16; i++) { int j = 1 + 4 * i; out[i] = (j < 16) ? tight[j] : 0u; }
    } else {
        NRC_REQUIRE(false, "unknown resolution mode");
    }
    return NRC_OK;
}

extern "C" nrc_status nrc_mrhe_create(nrc_ctx* ctx, float lr, float wd, uint32_t L, uint32_t T, uint32_t F, uint32_t minRes, uint32_t maxRes,
                                      const uint32_t* resolutions, uint64_t seed, nrc_mrhe** out)
{
    NRC_REQUIRE(ctx && out, "null argument");
    NRC_REQUIRE(L >= 1 && L <= 16, "levelCount must be 1..16");
    NRC_REQUIRE(F == 2, "featureCount must be 2 (the reference's only configuration)");
    NRC_REQUIRE(T >= 2 && (((size_t)L * T * F) % 4) == 0, "hashTableSize must make L*T*F a multiple of 4");
    NRC_CUDA(cudaSetDevice(ctx->device));
    nrc_mrhe* m = new nrc_mrhe();
    m->ctx = ctx; m->lr = lr; m->wd = wd; m->L = L; m->T = T; m->F = F; m->minRes = minRes; m->maxRes = maxRes;
    if (resolutions) {
        for (int i = 0; i < 16; i++) m->res[i] = resolutions[i];
    } else {
        NRC_TRY(nrc_mrhe_resolutions(minRes, maxRes, L < 2 ? 2 : L, NRC_RES_INTENDED, m->res));
    }
    m->count = (size_t)L * T * F;
    NRC_CUDA(cudaMalloc(&m->tables, m->count * sizeof(float)));
    NRC_CUDA(cudaMalloc(&m->delta, m->count * sizeof(float)));
    NRC_CUDA(cudaMemsetAsync(m->tables, 0, m->count * sizeof(float), ctx->stream));
    NRC_CUDA(cudaMemsetAsync(m->delta, 0, m->count * sizeof(float), ctx->stream));   // zero deltas
    if (seed != 0) NRC_TRY(launch_init_normal(ctx, m->tables, m->count, 0.1f, seed));   // MRHE.cpp:132-135
    *out = m;
    return NRC_OK;
}

extern "C" nrc_status nrc_mrhe_destroy(nrc_mrhe* m)
{
    if (!m) return NRC_OK;
    cudaStreamSynchronize(m->ctx->stream);
    cudaFree(m->tables); cudaFree(m->delta);
    if (m->m1) cudaFree(m->m1);
    if (m->m2) cudaFree(m->m2);
    delete m;
    return NRC_OK;
}

extern "C" nrc_status nrc_mrhe_set_config(nrc_mrhe* m, float lr, float wd)
{
    NRC_REQUIRE(m, "null mrhe");
    m->lr = lr; m->wd = wd;
    m->tailClean = false;
    return NRC_OK;
}

extern "C" size_t nrc_mrhe_table_bytes(nrc_mrhe* m) { return m ? m->count * sizeof(float) : 0; }

extern "C" nrc_status nrc_mrhe_set_tables(nrc_mrhe* m, const float* tables, const float* delta)
{
    NRC_REQUIRE(m, "null mrhe");
    if (tables) NRC_TRY(copy_in(m->ctx, m->tables, tables, m->count * sizeof(float)));
    if (delta) NRC_TRY(copy_in(m->ctx, m->delta, delta, m->count * sizeof(float)));
    m->tailClean = false;   // the next step must visit every entry again (steps.cu)
    return NRC_OK;
}

extern "C" nrc_status nrc_mrhe_get_tables(nrc_mrhe* m, float* tables, float* delta)
{
    NRC_REQUIRE(m, "null mrhe");
    if (tables) NRC_TRY(copy_out(m->ctx, tables, m->tables, m->count * sizeof(float)));
    if (delta) NRC_TRY(copy_out(m->ctx, delta, m->delta, m->count * sizeof(float)));
    return NRC_OK;
}

extern "C" float* nrc_mrhe_grad_ptr(nrc_mrhe* m) { return m ? m->delta : nullptr; }
extern "C" size_t nrc_mrhe_grad_extent(nrc_mrhe* m)
{
    if (!m) return 0;
    if (m->fixed) return m->count;   // the fixed preset scatters into every level's table
    size_t e = (size_t)2 * m->T;
    return e < m->count ? e : m->count;
}

// ---- the opt-in "fixed" preset ------------------------------------------------------------
extern "C" nrc_status nrc_mrhe_set_fixed_mode(nrc_mrhe* m, int on)
{
    NRC_REQUIRE(m, "null mrhe");
    m->fixed = on != 0;
    m->tailClean = false;
    return NRC_OK;
}

extern "C" nrc_status nrc_mlp_set_adam(nrc_mlp* mlp, int on, float beta2, float eps)
{
    NRC_REQUIRE(mlp, "null mlp");
    NRC_REQUIRE(!on || (beta2 > 0.0f && beta2 < 1.0f && eps >= 0.0f), "beta2 must be in (0, 1), eps >= 0");
    NRC_CUDA(cudaStreamSynchronize(mlp->ctx->stream));
    if (on) {
        if (!mlp->second) NRC_CUDA(cudaMalloc(&mlp->second, 21504 * sizeof(float)));
        NRC_CUDA(cudaMemset(mlp->second, 0, 21504 * sizeof(float)));
        mlp->beta2 = beta2; mlp->eps = eps;
    }
    mlp->adam = on != 0;
    mlp->steps = 0;
    return NRC_OK;
}

extern "C" nrc_status nrc_mrhe_set_adam(nrc_mrhe* m, int on, float beta1, float beta2, float eps)
{
    NRC_REQUIRE(m, "null mrhe");
    NRC_REQUIRE(!on || (beta1 >= 0.0f && beta1 < 1.0f && beta2 > 0.0f && beta2 < 1.0f && eps >= 0.0f), "betas must be in [0, 1), eps >= 0");
    NRC_CUDA(cudaStreamSynchronize(m->ctx->stream));
    if (on) {
        if (!m->m1) NRC_CUDA(cudaMalloc(&m->m1, m->count * sizeof(float)));
        if (!m->m2) NRC_CUDA(cudaMalloc(&m->m2, m->count * sizeof(float)));
        NRC_CUDA(cudaMemset(m->m1, 0, m->count * sizeof(float)));
        NRC_CUDA(cudaMemset(m->m2, 0, m->count * sizeof(float)));
        m->beta1 = beta1; m->beta2 = beta2; m->eps = eps;
    }
    m->adam = on != 0;
    m->tailClean = false;
    m->steps = 0;
    return NRC_OK;
}

// Adam moments for checkpoints (the first moment of the MLP is the reference's momentum: nrc_mlp_get_params / nrc_mlp_set_state)
extern "C" nrc_status nrc_mlp_get_adam_state(nrc_mlp* mlp, float* second, uint32_t* steps)
{
    NRC_REQUIRE(mlp && mlp->adam && mlp->second, "Adam is not switched on for this MLP");
    if (second) NRC_TRY(copy_out(mlp->ctx, second, mlp->second, NRC_MLP_PARAM_COUNT * sizeof(float)));
    if (steps) *steps = mlp->steps;
    return NRC_OK;
}

extern "C" nrc_status nrc_mlp_set_adam_state(nrc_mlp* mlp, const float* second, uint32_t steps)
{
    NRC_REQUIRE(mlp && mlp->adam && mlp->second, "Adam is not switched on for this MLP");
    if (second) NRC_TRY(copy_in(mlp->ctx, mlp->second, second, NRC_MLP_PARAM_COUNT * sizeof(float)));
    mlp->steps = steps;
    return NRC_OK;
}

extern "C" nrc_status nrc_mrhe_get_adam_state(nrc_mrhe* m, float* m1, float* m2, uint32_t* steps)
{
    NRC_REQUIRE(m && m->adam && m->m1 && m->m2, "Adam is not switched on for these tables");
    if (m1) NRC_TRY(copy_out(m->ctx, m1, m->m1, m->count * sizeof(float)));
    if (m2) NRC_TRY(copy_out(m->ctx, m2, m->m2, m->count * sizeof(float)));
    if (steps) *steps = m->steps;
    return NRC_OK;
}

extern "C" nrc_status nrc_mrhe_set_adam_state(nrc_mrhe* m, const float* m1, const float* m2, uint32_t steps)
{
    NRC_REQUIRE(m && m->adam && m->m1 && m->m2, "Adam is not switched on for these tables");
    if (m1) NRC_TRY(copy_in(m->ctx, m->m1, m1, m->count * sizeof(float)));
    if (m2) NRC_TRY(copy_in(m->ctx, m->m2, m2, m->count * sizeof(float)));
    m->steps = steps;
    return NRC_OK;
}

// ---- stand-alone operators ------------------------------------------------------------------------------------
extern "C" nrc_status nrc_encode(nrc_mrhe* m, const float* pos, size_t n, float* feat, uint32_t* idx)
{
    NRC_REQUIRE(m && pos, "null argument");
    if (n == 0) return NRC_OK;
    nrc_ctx* ctx = m->ctx;
    const void* dpos; void* dfeat = nullptr; void* didx = nullptr;
    NRC_TRY(nrc_stage_in(ctx, 0, pos, n * 3 * sizeof(float), &dpos));
    if (feat) NRC_TRY(nrc_stage_out(ctx, 3, feat, n * 32 * sizeof(float), &dfeat));
    if (idx) NRC_TRY(nrc_stage_out(ctx, 4, idx, n * 128 * sizeof(uint32_t), &didx));
    NRC_TRY(launch_encode_pos(m, (const float*)dpos, n, (float*)dfeat, (uint32_t*)didx));
    if (feat) NRC_TRY(nrc_stage_out_finish(ctx, feat, dfeat, n * 32 * sizeof(float)));
    if (idx) NRC_TRY(nrc_stage_out_finish(ctx, idx, didx, n * 128 * sizeof(uint32_t)));
    return NRC_OK;
}

extern "C" nrc_status nrc_encode_dir(nrc_ctx* ctx, const float* dir, size_t n, float* feat)
{
    NRC_REQUIRE(ctx && dir && feat, "null argument");
    if (n == 0) return NRC_OK;
    const void* ddir; void* dfeat;
    NRC_TRY(nrc_stage_in(ctx, 0, dir, n * 3 * sizeof(float), &ddir));
    NRC_TRY(nrc_stage_out(ctx, 3, feat, n * 32 * sizeof(float), &dfeat));
    NRC_TRY(launch_encode_dir(ctx, (const float*)ddir, n, (float*)dfeat));
    NRC_TRY(nrc_stage_out_finish(ctx, feat, dfeat, n * 32 * sizeof(float)));
    return NRC_OK;
}

extern "C" nrc_status nrc_infer(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, size_t n, float* rgb)
{
    NRC_REQUIRE(mlp && m && pos && dir && rgb, "null argument");
    NRC_REQUIRE(mlp->ctx == m->ctx, "mlp and mrhe belong to different contexts");
    if (n == 0) return NRC_OK;
    nrc_ctx* ctx = mlp->ctx;
    const void *dpos, *ddir; void* drgb;
    NRC_TRY(nrc_stage_in(ctx, 0, pos, n * 3 * sizeof(float), &dpos));
    NRC_TRY(nrc_stage_in(ctx, 1, dir, n * 3 * sizeof(float), &ddir));
    NRC_TRY(nrc_stage_out(ctx, 3, rgb, n * 3 * sizeof(float), &drgb));
    if (ctx->precision == NRC_MLP_BF16_TC && ctx->inferMode == NRC_INFER_SPLIT)
        NRC_TRY(launch_infer_split(mlp, m, (const float*)dpos, (const float*)ddir, nullptr, n, (float*)drgb, ctx->stream));
    else if (ctx->precision == NRC_MLP_BF16_TC)
        NRC_TRY(launch_infer_tc(mlp, m, (const float*)dpos, (const float*)ddir, nullptr, n, (float*)drgb, ctx->stream));
    else
        NRC_TRY(launch_infer_fp32(mlp, m, (const float*)dpos, (const float*)ddir, nullptr, n, (float*)drgb, ctx->stream));
    NRC_TRY(nrc_stage_out_finish(ctx, rgb, drgb, n * 3 * sizeof(float)));
    return NRC_OK;
}

extern "C" size_t nrc_feature_tile_bytes(size_t n) { return ((n + 127) / 128) * 16384; }

extern "C" nrc_status nrc_encode_tiles(nrc_mrhe* m, const float* pos, const float* dir, size_t n, void* featTiles)
{
    NRC_REQUIRE(m && pos && dir && featTiles, "null argument");
    NRC_REQUIRE(nrc_is_device_ptr(featTiles), "featTiles must be device memory");
    if (n == 0) return NRC_OK;
    nrc_ctx* ctx = m->ctx;
    const void *dpos, *ddir;
    NRC_TRY(nrc_stage_in(ctx, 0, pos, n * 3 * sizeof(float), &dpos));
    NRC_TRY(nrc_stage_in(ctx, 1, dir, n * 3 * sizeof(float), &ddir));
    return launch_encode_tiles(m, (const float*)dpos, (const float*)ddir, nullptr, n, featTiles, ctx->stream);
}

extern "C" nrc_status nrc_mlp_forward_tiles(nrc_mlp* mlp, const void* featTiles, size_t n, float* rgb)
{
    NRC_REQUIRE(mlp && featTiles && rgb, "null argument");
    NRC_REQUIRE(nrc_is_device_ptr(featTiles), "featTiles must be device memory");
    if (n == 0) return NRC_OK;
    nrc_ctx* ctx = mlp->ctx;
    void* drgb;
    NRC_TRY(nrc_stage_out(ctx, 3, rgb, n * 3 * sizeof(float), &drgb));
    NRC_TRY(launch_mlp_stream(mlp, featTiles, nullptr, n, (float*)drgb, ctx->stream));
    NRC_TRY(nrc_stage_out_finish(ctx, rgb, drgb, n * 3 * sizeof(float)));
    return NRC_OK;
}

extern "C" nrc_status nrc_train_accumulate(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const float* target, size_t n,
                                           float lossScale, float* pred)
{
    NRC_REQUIRE(mlp && m && pos && dir && target, "null argument");
    NRC_REQUIRE(mlp->ctx == m->ctx, "mlp and mrhe belong to different contexts");
    if (n == 0) return NRC_OK;
    nrc_ctx* ctx = mlp->ctx;
    const void *dpos, *ddir, *dtgt; void* dpred = nullptr;
    NRC_TRY(nrc_stage_in(ctx, 0, pos, n * 3 * sizeof(float), &dpos));
    NRC_TRY(nrc_stage_in(ctx, 1, dir, n * 3 * sizeof(float), &ddir));
    NRC_TRY(nrc_stage_in(ctx, 2, target, n * 3 * sizeof(float), &dtgt));
    if (pred) NRC_TRY(nrc_stage_out(ctx, 3, pred, n * 3 * sizeof(float), &dpred));
    if (ctx->precision == NRC_MLP_BF16_TC)
        NRC_TRY(launch_train_tc(mlp, m, (const float*)dpos, (const float*)ddir, (const float*)dtgt, n, lossScale, (float*)dpred, ctx->stream));
    else
        NRC_TRY(launch_train_fp32(mlp, m, (const float*)dpos, (const float*)ddir, (const float*)dtgt, n, lossScale, (float*)dpred, ctx->stream));
    if (pred) NRC_TRY(nrc_stage_out_finish(ctx, pred, dpred, n * 3 * sizeof(float)));
    return NRC_OK;
}

extern "C" nrc_status nrc_optimizer_step(nrc_mlp* mlp, nrc_mrhe* m)
{
    NRC_REQUIRE(mlp || m, "nothing to step");
    if (mlp) NRC_TRY(launch_nrc_step(mlp, mlp->ctx->stream));
    if (m) NRC_TRY(launch_mrhe_step(m, m->ctx->stream));
    return NRC_OK;
}
