// renderer.cu -- nrc_renderer: the frame graph that replaces NrcHpmRenderer's recorded command buffer
// (src/NrcHpmRenderer.cpp:571-730): train -> nrc-step -> mrhe-step -> render.
//
//   walk streams : train walk            } independent of the weights AND of each other: consecutive frames' walks run on alternating
//                                        } streams into NRC_TRAIN_SETS buffer sets, so up to three of these latency-bound chains are in flight
//   stream4 : train light -> combine     } at once; all of it overlaps frame k's steps / exchange wait / inference / composite
//   stream  : fused encode+MLP fwd/bwd -> [gradient exchange] -> steps -> batched inference -> composite
//   stream2 : render walk -> light -> combine (compacts the NRC queries)           (independent of the weights)
#include <math.h>
#include <string.h>
#include <stdlib.h>
#include "trace.cuh"

struct nrc_renderer {
    nrc_ctx* ctx = nullptr;
    nrc_volume* vol = nullptr;
    nrc_envmap* env = nullptr;
    nrc_mlp* mlp = nullptr;
    nrc_mrhe* mrhe = nullptr;
    uint32_t width = 0, height = 0, trainW = 0, trainH = 0;
    uint32_t y0 = 0, y1 = 0, i0 = 0, i1 = 0;   // this rank's partition: rows y0, y0+ystride, ... < y1; invocations i0, i0+istride, ... < i1
    uint32_t ystride = 1, istride = 1;
    nrc_camera cam;
    nrc_dir_light dirLight;
    nrc_point_light ptLight;
    // frame
    float4* frame = nullptr;
    uint32_t* frame8 = nullptr;
    float4* background = nullptr;
    uint32_t* counters = nullptr;     // [h*w][4]
    float* queryProbe = nullptr;      // [h*w][8]
    // render paths / pool / queries: NRC_RENDER_SETS sets, so that the trace of frame k+1 (and the walk of k+2) never waits for frame k's inference + composite.
    // rp / rpool / ql / background alias the set of the last nrc_frame_render (what the probes read).
    uint32_t sppAlloc = 0;
    RenderPaths rp;
    VertexPool rpool;
    QueryList ql;
    struct RenderSet { RenderPaths rp; VertexPool rpool; QueryList ql; float4* background; } rs[NRC_RENDER_SETS];
    int renderBuf = 1;
    bool setPending[NRC_RENDER_SETS] = {};  // evSetFree[b]: the composite that last read set b has finished
    cudaEvent_t evSetFree[NRC_RENDER_SETS] = {};
    cudaEvent_t evScalars = nullptr;      // the last frame's device scalars have reached hostScalars
    bool overflowPending = false;         // the last frame used a pool below the worst case and its overflow flag is unread
    TraceParams lastP;                    // what the last frame was traced with (a redo after an overflow repeats it)
    uint32_t lastSpp = 0;
    cudaEvent_t evCopyReady = nullptr, evCopyDone[4] = {nullptr, nullptr, nullptr, nullptr};   // asynchronous read-back
    int copyNext = 0, copyLast = -1;
    size_t poolCapHint = 0;               // vertex-pool capacity with the cache on, set by the caller (0: the library chooses, frame_render)
    size_t autoPaths = 0;                 // the library's choice was made for this many paths ...
    bool autoWorst = false;               // ... and is the worst-case pool
    bool lastInfer = false;
    // train
    TrainPaths tp;                    // = tpb[trainBuf]: the buffers of the last nrc_frame_train (what the probes read)
    VertexPool tpool;
    TrainPaths tpb[NRC_TRAIN_SETS];   // multi-buffered: the walks of frames k+1.. fill other sets while lights/combine/MLP of frame k read theirs
    VertexPool tpoolb[NRC_TRAIN_SETS];
    int trainBuf = NRC_TRAIN_SETS - 1;
    unsigned trainSeq = 0;            // train frames queued so far: picks the walk stream
    bool trainEnvLight = false;       // opt-in: env-map term in the training targets (train pools then carry random numbers)
    uint32_t* trainCounters = nullptr;             // = trainCountersB[trainBuf]
    uint32_t* trainCountersB[NRC_TRAIN_SETS] = {nullptr, nullptr, nullptr, nullptr};   // per set: walks of different frames run concurrently
    uint32_t trainAlloc = 0;
    // device scalars, render set b at base 8*b: [0] vertex count, [1] overflow, [4] query count, [5] render-walk paths handed to the tail launch;
    // train set b: {32 + 2b, 33 + 2b} = vertex count, overflow
    uint32_t* scalars = nullptr;
    unsigned long long* taps = nullptr;   // [3 b] render taps of set b; [12 + b] train taps of set b
    unsigned long long* totals = nullptr; // [6] running totals (nrc_frame_totals)
    uint32_t* hostScalars = nullptr;      // pinned mirror
    unsigned long long* hostTaps = nullptr;
    bool trainPending[NRC_TRAIN_SETS] = {false, false, false, false};  // evTrainFree[b] marks the end of the train kernel that last consumed buffer set b
};

#define FREE(p) do { if (p) { cudaFree(p); (p) = nullptr; } } while (0)

static void free_pool(VertexPool& p) { FREE(p.a); FREE(p.b); FREE(p.c); FREE(p.rnd); FREE(p.out); p.cap = 0; }

static nrc_status alloc_pool(VertexPool& p, uint32_t cap, bool withRnd, uint32_t* count, uint32_t* overflow)
{
    free_pool(p);
    NRC_CUDA(cudaMalloc(&p.a, (size_t)cap * sizeof(float4)));
    NRC_CUDA(cudaMalloc(&p.b, (size_t)cap * sizeof(float4)));
    NRC_CUDA(cudaMalloc(&p.c, (size_t)cap * sizeof(float4)));
    NRC_CUDA(cudaMalloc(&p.out, (size_t)cap * sizeof(float4)));
    if (withRnd) NRC_CUDA(cudaMalloc(&p.rnd, (size_t)cap * 4 * sizeof(float4)));
    p.cap = cap; p.count = count; p.overflow = overflow;
    return NRC_OK;
}

static void free_frame(nrc_renderer* r)
{
    FREE(r->frame); FREE(r->frame8); FREE(r->counters); FREE(r->queryProbe);
    for (int b = 0; b < NRC_RENDER_SETS; b++) {
        nrc_renderer::RenderSet& S = r->rs[b];
        FREE(S.background);
        FREE(S.rp.head); FREE(S.rp.qposPhase); FREE(S.rp.qdirValid); FREE(S.rp.result); FREE(S.rp.qweight); FREE(S.rp.qslot); FREE(S.rp.spill);
        FREE(S.ql.pos); FREE(S.ql.dir); FREE(S.ql.rgb);
        free_pool(S.rpool);
        r->setPending[b] = false;
    }
    memset(&r->rp, 0, sizeof(r->rp)); memset(&r->rpool, 0, sizeof(r->rpool)); memset(&r->ql, 0, sizeof(r->ql));
    r->background = nullptr;
    r->overflowPending = false;
    r->sppAlloc = 0;
}

static void use_render_set(nrc_renderer* r, int b)
{
    r->renderBuf = b;
    r->rp = r->rs[b].rp; r->rpool = r->rs[b].rpool; r->ql = r->rs[b].ql; r->background = r->rs[b].background;
}

// (Re)allocate everything whose size depends on the frame size and spp.  The caller has drained every stream.
static nrc_status alloc_frame(nrc_renderer* r, uint32_t spp, uint32_t poolCap)
{
    free_frame(r);
    size_t px = (size_t)r->width * r->height;
    size_t paths = px * spp;
    NRC_CUDA(cudaMalloc(&r->frame, px * sizeof(float4)));
    NRC_CUDA(cudaMalloc(&r->frame8, px * sizeof(uint32_t)));
    NRC_CUDA(cudaMalloc(&r->counters, px * 4 * sizeof(uint32_t)));
    NRC_CUDA(cudaMalloc(&r->queryProbe, px * 8 * sizeof(float)));
    NRC_CUDA(cudaMemsetAsync(r->frame, 0, px * sizeof(float4), r->ctx->stream));
    NRC_CUDA(cudaMemsetAsync(r->frame8, 0, px * sizeof(uint32_t), r->ctx->stream));
    NRC_CUDA(cudaMemsetAsync(r->counters, 0, px * 4 * sizeof(uint32_t), r->ctx->stream));
    NRC_CUDA(cudaMemsetAsync(r->queryProbe, 0, px * 8 * sizeof(float), r->ctx->stream));
    for (int b = 0; b < NRC_RENDER_SETS; b++) {
        nrc_renderer::RenderSet& S = r->rs[b];
        NRC_CUDA(cudaMalloc(&S.background, px * sizeof(float4)));
        NRC_CUDA(cudaMalloc(&S.rp.head, paths * sizeof(int)));
        NRC_CUDA(cudaMalloc(&S.rp.qposPhase, paths * sizeof(float4)));
        NRC_CUDA(cudaMalloc(&S.rp.qdirValid, paths * sizeof(float4)));
        NRC_CUDA(cudaMalloc(&S.rp.result, paths * sizeof(float4)));
        NRC_CUDA(cudaMalloc(&S.rp.qweight, paths * sizeof(float)));
        NRC_CUDA(cudaMalloc(&S.rp.qslot, paths * sizeof(int)));
        S.rp.spill = nullptr;                      // state of the paths the head launch of the render walk leaves alive (trace.cu)
        S.rp.spillStride = 0;
        S.rp.spillCount = r->scalars + 8 * b + 5;
        if (spp == 1) {
            NRC_CUDA(cudaMalloc(&S.rp.spill, px * 5 * sizeof(float4)));
            S.rp.spillStride = (uint32_t)px;
        }
        NRC_CUDA(cudaMalloc(&S.ql.pos, paths * 3 * sizeof(float)));
        NRC_CUDA(cudaMalloc(&S.ql.dir, paths * 3 * sizeof(float)));
        NRC_CUDA(cudaMalloc(&S.ql.rgb, paths * 3 * sizeof(float)));
        S.ql.cap = (uint32_t)paths;
        S.ql.count = r->scalars + 8 * b + 4;
        NRC_TRY(alloc_pool(S.rpool, poolCap, true, r->scalars + 8 * b + 0, r->scalars + 8 * b + 1));
    }
    use_render_set(r, 1);
    r->sppAlloc = spp;
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream));
    return NRC_OK;
}

static void free_train(nrc_renderer* r)
{
    for (int b = 0; b < NRC_TRAIN_SETS; b++) {
        FREE(r->tpb[b].head); FREE(r->tpb[b].pos); FREE(r->tpb[b].dir); FREE(r->tpb[b].target);
        free_pool(r->tpoolb[b]);
        FREE(r->trainCountersB[b]);
    }
    r->trainCounters = nullptr;
    memset(&r->tp, 0, sizeof(r->tp)); memset(&r->tpool, 0, sizeof(r->tpool));
    r->trainAlloc = 0;
}

static nrc_status alloc_train(nrc_renderer* r)
{
    free_train(r);
    uint32_t n = r->trainW * r->trainH;
    size_t cap = (size_t)n * (8 * 64 + 2 * 16);   // worst case: 8 paths x 64 lit vertices per training sample (+ the two partly
                                                  // used chunks a thread can end with, TRAIN_CHUNK) -- this pool cannot overflow
    NRC_REQUIRE(cap < ((size_t)1 << 31), "train grid too large for the vertex pool");
    for (int b = 0; b < NRC_TRAIN_SETS; b++) {
        TrainPaths& tp = r->tpb[b];
        NRC_CUDA(cudaMalloc(&tp.head, (size_t)n * 8 * sizeof(int)));
        NRC_CUDA(cudaMalloc(&tp.pos, (size_t)n * 3 * sizeof(float)));
        NRC_CUDA(cudaMalloc(&tp.dir, (size_t)n * 3 * sizeof(float)));
        NRC_CUDA(cudaMalloc(&tp.target, (size_t)n * 3 * sizeof(float)));
        NRC_CUDA(cudaMemsetAsync(tp.pos, 0, (size_t)n * 3 * sizeof(float), r->ctx->stream));
        NRC_CUDA(cudaMemsetAsync(tp.dir, 0, (size_t)n * 3 * sizeof(float), r->ctx->stream));
        NRC_CUDA(cudaMemsetAsync(tp.target, 0, (size_t)n * 3 * sizeof(float), r->ctx->stream));
        NRC_TRY(alloc_pool(r->tpoolb[b], (uint32_t)cap, r->trainEnvLight, r->scalars + 32 + 2 * b, r->scalars + 33 + 2 * b));
    }
    for (int b = 0; b < NRC_TRAIN_SETS; b++) {
        NRC_CUDA(cudaMalloc(&r->trainCountersB[b], (size_t)n * 4 * sizeof(uint32_t)));
        NRC_CUDA(cudaMemsetAsync(r->trainCountersB[b], 0, (size_t)n * 4 * sizeof(uint32_t), r->ctx->stream));
    }
    r->trainBuf = NRC_TRAIN_SETS - 1;
    r->tp = r->tpb[r->trainBuf]; r->tpool = r->tpoolb[r->trainBuf]; r->trainCounters = r->trainCountersB[r->trainBuf];
    r->trainAlloc = n;
    return NRC_OK;
}

// ---- create / destroy / setters ------------------------------------------------------------------------------------------
extern "C" nrc_status nrc_renderer_create(nrc_ctx* ctx, uint32_t width, uint32_t height, uint32_t trainWidth, uint32_t trainHeight,
                                          nrc_volume* vol, nrc_envmap* env, nrc_mlp* mlp, nrc_mrhe* mrhe, nrc_renderer** out)
{
    NRC_REQUIRE(ctx && vol && env && mlp && mrhe && out, "null argument");
    NRC_REQUIRE(width > 0 && height > 0 && trainWidth > 0 && trainHeight > 0, "empty frame or train grid");
    NRC_REQUIRE((size_t)width * height < ((size_t)1 << 28), "frame too large");
    NRC_CUDA(cudaSetDevice(ctx->device));
    nrc_renderer* r = new nrc_renderer();
    r->ctx = ctx; r->vol = vol; r->env = env; r->mlp = mlp; r->mrhe = mrhe;
    r->width = width; r->height = height; r->trainW = trainWidth; r->trainH = trainHeight;
    r->y0 = 0; r->y1 = height; r->i0 = 0; r->i1 = trainWidth * trainHeight;
    memset(&r->cam, 0, sizeof(r->cam));
    memset(&r->dirLight, 0, sizeof(r->dirLight));
    memset(&r->ptLight, 0, sizeof(r->ptLight));
    memset(&r->rp, 0, sizeof(r->rp)); memset(&r->rpool, 0, sizeof(r->rpool)); memset(&r->ql, 0, sizeof(r->ql));
    memset(&r->tp, 0, sizeof(r->tp)); memset(&r->tpool, 0, sizeof(r->tpool));
    memset(r->tpb, 0, sizeof(r->tpb)); memset(r->tpoolb, 0, sizeof(r->tpoolb));
    NRC_CUDA(cudaMalloc(&r->scalars, 64 * sizeof(uint32_t)));
    NRC_CUDA(cudaMalloc(&r->taps, 16 * sizeof(unsigned long long)));
    NRC_CUDA(cudaMalloc(&r->totals, 6 * sizeof(unsigned long long)));
    NRC_CUDA(cudaMemsetAsync(r->totals, 0, 6 * sizeof(unsigned long long), ctx->stream));
    NRC_CUDA(cudaMemsetAsync(r->scalars, 0, 64 * sizeof(uint32_t), ctx->stream));
    memset(r->rs, 0, sizeof(r->rs));
    for (int b = 0; b < NRC_RENDER_SETS; b++) NRC_CUDA(cudaEventCreateWithFlags(&r->evSetFree[b], cudaEventDisableTiming));
    NRC_CUDA(cudaEventCreateWithFlags(&r->evScalars, cudaEventDisableTiming));
    NRC_CUDA(cudaEventCreateWithFlags(&r->evCopyReady, cudaEventDisableTiming));
    for (int k = 0; k < 4; k++) NRC_CUDA(cudaEventCreateWithFlags(&r->evCopyDone[k], cudaEventDisableTiming));
    NRC_CUDA(cudaMemsetAsync(r->taps, 0, 16 * sizeof(unsigned long long), ctx->stream));
    NRC_CUDA(cudaMallocHost(&r->hostScalars, 8 * sizeof(uint32_t)));
    NRC_CUDA(cudaMallocHost(&r->hostTaps, 16 * sizeof(unsigned long long)));
    memset(r->hostScalars, 0, 8 * sizeof(uint32_t));
    memset(r->hostTaps, 0, 16 * sizeof(unsigned long long));
    NRC_TRY(alloc_frame(r, 1, (uint32_t)((size_t)width * height * 4)));
    NRC_TRY(alloc_train(r));
    ctx->renderers.push_back(r);
    *out = r;
    return NRC_OK;
}

extern "C" nrc_status nrc_renderer_destroy(nrc_renderer* r)
{
    if (!r) return NRC_OK;
    for (size_t i = 0; i < r->ctx->renderers.size(); i++)
        if (r->ctx->renderers[i] == r) { r->ctx->renderers.erase(r->ctx->renderers.begin() + i); break; }
    cudaStreamSynchronize(r->ctx->stream);
    cudaStreamSynchronize(r->ctx->stream2);
    for (int i = 0; i < NRC_TRAIN_WALK_STREAMS; i++) cudaStreamSynchronize(r->ctx->streamWalk[i]);
    cudaStreamSynchronize(r->ctx->stream4);
    cudaStreamSynchronize(r->ctx->stream5);
    cudaStreamSynchronize(r->ctx->streamCopy);
    free_frame(r);
    free_train(r);
    FREE(r->scalars); FREE(r->taps); FREE(r->totals);
    for (int b = 0; b < NRC_RENDER_SETS; b++) if (r->evSetFree[b]) cudaEventDestroy(r->evSetFree[b]);
    if (r->evScalars) cudaEventDestroy(r->evScalars);
    if (r->evCopyReady) cudaEventDestroy(r->evCopyReady);
    for (int k = 0; k < 4; k++) if (r->evCopyDone[k]) cudaEventDestroy(r->evCopyDone[k]);
    if (r->hostScalars) cudaFreeHost(r->hostScalars);
    if (r->hostTaps) cudaFreeHost(r->hostTaps);
    delete r;
    return NRC_OK;
}

extern "C" nrc_status nrc_renderer_set_camera(nrc_renderer* r, const nrc_camera* cam)
{
    NRC_REQUIRE(r && cam, "null argument");
    r->cam = *cam;
    return NRC_OK;
}

extern "C" nrc_status nrc_renderer_set_dir_light(nrc_renderer* r, const nrc_dir_light* l)
{
    NRC_REQUIRE(r && l, "null argument");
    r->dirLight = *l;
    return NRC_OK;
}

extern "C" nrc_status nrc_renderer_set_point_light(nrc_renderer* r, const nrc_point_light* l)
{
    NRC_REQUIRE(r && l, "null argument");
    r->ptLight = *l;
    return NRC_OK;
}

extern "C" nrc_status nrc_renderer_set_partition(nrc_renderer* r, uint32_t y0, uint32_t y1, uint32_t i0, uint32_t i1)
{
    NRC_REQUIRE(r, "null renderer");
    NRC_REQUIRE(y0 <= y1 && y1 <= r->height, "row partition out of range");
    NRC_REQUIRE(i0 <= i1 && i1 <= r->trainW * r->trainH, "train partition out of range");
    r->y0 = y0; r->y1 = y1; r->i0 = i0; r->i1 = i1;
    r->ystride = 1; r->istride = 1;
    return NRC_OK;
}

extern "C" nrc_status nrc_renderer_set_partition_strided(nrc_renderer* r, uint32_t yFirst, uint32_t yStride, uint32_t iFirst, uint32_t iStride)
{
    NRC_REQUIRE(r, "null renderer");
    NRC_REQUIRE(yStride >= 1 && iStride >= 1 && yFirst < yStride && iFirst < iStride, "bad strided partition");
    r->y0 = yFirst; r->y1 = r->height; r->ystride = yStride;
    r->i0 = iFirst; r->i1 = r->trainW * r->trainH; r->istride = iStride;
    return NRC_OK;
}

static nrc_status resolve_overflow_fwd(nrc_renderer* r);
static inline uint32_t part_count(uint32_t first, uint32_t end, uint32_t stride) { return first >= end ? 0u : (end - first + stride - 1) / stride; }

extern "C" nrc_status nrc_resize(nrc_renderer* r, uint32_t width, uint32_t height)
{
    NRC_REQUIRE(r && width > 0 && height > 0, "bad size");
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream2));
    for (int i = 0; i < NRC_TRAIN_WALK_STREAMS; i++) NRC_CUDA(cudaStreamSynchronize(r->ctx->streamWalk[i]));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream4));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream5));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->streamCopy));
    r->width = width; r->height = height;
    r->y0 = r->ystride > 1 ? r->y0 : 0; r->y1 = height;
    return alloc_frame(r, r->sppAlloc ? r->sppAlloc : 1, (uint32_t)((size_t)width * height * 4));
}

// Initial capacity of the render vertex pool while the cache is on (0 = the default of 4 vertices per path; 32 per path
// cannot overflow).  A frame that overflows a smaller pool is traced again with the worst-case pool before its results are
// handed out.
extern "C" nrc_status nrc_renderer_set_pool_capacity(nrc_renderer* r, uint64_t vertices)
{
    NRC_REQUIRE(r, "null renderer");
    NRC_REQUIRE(vertices < 0x7fffffffull, "pool capacity out of range");
    NRC_TRY(resolve_overflow_fwd(r));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream2));
    for (int i = 0; i < NRC_TRAIN_WALK_STREAMS; i++) NRC_CUDA(cudaStreamSynchronize(r->ctx->streamWalk[i]));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream4));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream5));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->streamCopy));
    r->poolCapHint = (size_t)vertices;
    uint32_t spp = r->sppAlloc ? r->sppAlloc : 1;
    size_t cap = vertices ? (size_t)vertices : (size_t)r->width * r->height * spp * 4;
    if (cap > 0x7fffffffu) cap = 0x7fffffffu;
    return alloc_frame(r, spp, (uint32_t)cap);
}

// Opt-in estimator: the env-map term of TraceScene in the TRAINING targets, which nrc-train.comp:1118 has
// commented out.  A lit training vertex then draws the 16 random numbers of SampleHdrEnvMap first (the render walk's order),
// the train pools carry them, and the training vertices go through the render variant of the light kernel.
extern "C" nrc_status nrc_renderer_set_train_env_light(nrc_renderer* r, int on)
{
    NRC_REQUIRE(r, "null renderer");
    if (r->trainEnvLight == (on != 0)) return NRC_OK;
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream2));
    for (int i = 0; i < NRC_TRAIN_WALK_STREAMS; i++) NRC_CUDA(cudaStreamSynchronize(r->ctx->streamWalk[i]));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream4));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream5));
    r->trainEnvLight = on != 0;
    for (int b = 0; b < NRC_TRAIN_SETS; b++) r->trainPending[b] = false;
    return alloc_train(r);
}

static TraceParams make_params(const nrc_renderer* r)
{
    TraceParams P;
    memset(&P, 0, sizeof(P));
    memcpy(P.invProjView, r->cam.invProjView, sizeof(P.invProjView));
    memcpy(P.camPos, r->cam.pos, sizeof(P.camPos));
    const nrc_volume_params& v = r->vol->params;
    memcpy(P.random, v.random, sizeof(P.random));
    P.useNN = v.useNN; P.showNonNN = v.showNonNN; P.densityFactor = v.densityFactor; P.g = v.g;
    P.noNnSpp = v.noNnSpp; P.withNnSpp = v.withNnSpp;
    memcpy(P.dirDir, r->dirLight.dir, sizeof(P.dirDir));
    P.dirStrength = r->dirLight.strength;
    memcpy(P.ptPos, r->ptLight.pos, sizeof(P.ptPos));
    P.ptStrength = r->ptLight.strength;
    memcpy(P.ptColor, r->ptLight.color, sizeof(P.ptColor));
    P.envDirect = r->env->directStrength; P.envHpm = r->env->hpmStrength;
    P.skySize[0] = 62.5f; P.skySize[1] = 42.5f; P.skySize[2] = 76.5f;   // vec3(125,85,153)/2 (nrc-train.comp:212)
    // length(2*skySize) with the oracle's evaluation order: sqrt((x*x + y*y) + z*z), IEEE float on the host
    {
        volatile float x = 2.0f * P.skySize[0], y = 2.0f * P.skySize[1], z = 2.0f * P.skySize[2];
        volatile float xx = x * x, yy = y * y, zz = z * z;
        volatile float s = xx + yy;
        volatile float t = s + zz;
        P.exitJump = sqrtf(t);
    }
    P.nx = (int)r->vol->nx; P.ny = (int)r->vol->ny; P.nz = (int)r->vol->nz;
    P.bricks = r->vol->bricks;
    P.env = r->env->rgba; P.envW = (int)r->env->w; P.envH = (int)r->env->h;
    const bool cdfHalf = r->env->cdfHalf && r->env->invCdfX && r->env->invCdfY;
    P.envInvCdfX = cdfHalf ? r->env->invCdfX : nullptr; P.envInvCdfY = cdfHalf ? r->env->invCdfY : nullptr;
    P.width = r->width; P.height = r->height; P.trainW = r->trainW; P.trainH = r->trainH;
    P.widthFactor = 1.0f / (float)r->trainW;     // NrcHpmRenderer.cpp:409-411
    P.heightFactor = 1.0f / (float)r->trainH;
    return P;
}

// ---- the frame ---------------------------------------------------------------------------------------------------------------
static nrc_status infer_queries(nrc_renderer* r, cudaStream_t st)
{
    if (r->ctx->precision == NRC_MLP_BF16_TC)
        return (r->ctx->inferMode == NRC_INFER_SPLIT ? launch_infer_split : launch_infer_tc)(r->mlp, r->mrhe, r->ql.pos, r->ql.dir, r->ql.count,
                                                                                             r->ql.cap, r->ql.rgb, st);
    return launch_infer_fp32(r->mlp, r->mrhe, r->ql.pos, r->ql.dir, r->ql.count, r->ql.cap, r->ql.rgb, st);
}

// nrc-train.comp over this rank's invocations: trace the records, then Backprop() them (gradients accumulate
// into the MLP / MRHE delta arenas; the optimiser step is separate).
// deferred: see launch_train_* -- nrc_render finishes the frame's training with one fused reduce + steps launch
static nrc_status frame_train(nrc_renderer* r, int* deferred)
{
    NRC_REQUIRE(r, "null renderer");
    nrc_ctx* ctx = r->ctx;
    cudaStream_t st = ctx->stream;
    uint32_t n = part_count(r->i0, r->i1, r->istride);
    if (n == 0) return NRC_OK;
    TraceParams P = make_params(r);
    // The trace half (walk, lights, combine) reads no weights, so it runs AHEAD of the main stream, on two streams of its own:
    //   walk streams: walk(k)                  -- the long pole (one serial chain per sample); walk(k+1), walk(k+2) run beside it on the
    //                                             other walk streams (NRC_TRAIN_WALK=n: walks in flight, default 3; 1 = one after the other)
    //   stream4: light(k) -> combine(k)        -- overlaps the later walks, which fill other buffer sets
    //   main   : train kernel(k) -> [exchange] -> steps -> inference -> composite
    // so frame k+1's trace overlaps frame k's steps / inference / composite and, on several GPUs, the exchange's wait for
    // the slowest rank.  A buffer set is reused once the train kernel that read it has finished (evTrainFree[b]).
    static int noOverlap = -1, noAhead = -1;
    if (noOverlap < 0) { const char* e = getenv("NRC_NO_OVERLAP"); noOverlap = (e && atoi(e)) ? 1 : 0; }
    if (noAhead < 0) { const char* e = getenv("NRC_NO_TRAIN_AHEAD"); noAhead = (e && atoi(e)) ? 1 : 0; }
    const bool ahead = !(noOverlap || noAhead);
    // Walks in flight: three (all walk streams).  The walk is one dependent chain per sample (~1.2 ms alone, ~2.5 ms beside the other
    // kernels); with the frames' walks overlapping it never bounds the frame: 720p 1.80 against 2.15 ms with one at a time, 1080p 2.77
    // against 2.81 ms, 4K per rank the same either way (profiles/r02_walks_in_flight.md, r02_carveout.md).  NRC_TRAIN_WALK=n overrides.
    static int walkEnv = -1;
    if (walkEnv < 0) {
        const char* e = getenv("NRC_TRAIN_WALK");
        walkEnv = e ? atoi(e) : 0;
        if (walkEnv < 1 || walkEnv > NRC_TRAIN_WALK_STREAMS) walkEnv = NRC_TRAIN_WALK_STREAMS;
    }
    const int walkStreams = walkEnv;
    cudaStream_t tw = ahead ? ctx->streamWalk[r->trainSeq % (unsigned)walkStreams] : st;
    cudaStream_t tl = ahead ? ctx->stream4 : st;
    r->trainSeq++;
    const int b = ahead ? (r->trainBuf + 1) % NRC_TRAIN_SETS : r->trainBuf;
    r->trainBuf = b;
    r->tp = r->tpb[b]; r->tpool = r->tpoolb[b]; r->trainCounters = r->trainCountersB[b];
    if (ahead && r->trainPending[b]) NRC_CUDA(cudaStreamWaitEvent(tw, ctx->evTrainFree[b], 0));
    NRC_CUDA(cudaMemsetAsync(r->tpool.count, 0, 2 * sizeof(uint32_t), tw));          // count + overflow of this set
    NRC_CUDA(cudaMemsetAsync(r->taps + 12 + b, 0, sizeof(unsigned long long), tw));
    NRC_TRY(launch_train_walk(ctx, tw, P, r->i0, n, r->istride, r->tpool, r->tp, r->trainCounters));
    if (ahead) {
        NRC_CUDA(cudaEventRecord(ctx->evTrainWalk, tw));
        NRC_CUDA(cudaStreamWaitEvent(tl, ctx->evTrainWalk, 0));
    }
    NRC_TRY(launch_light(ctx, tl, P, r->tpool, r->trainEnvLight, r->taps + 12 + b, K_LIGHT_TRAIN));
    NRC_TRY(launch_train_combine(ctx, tl, n, r->tpool, r->tp, r->taps + 12 + b, r->totals));
    if (ahead) {
        NRC_CUDA(cudaEventRecord(ctx->evTrainTrace, tl));
        NRC_CUDA(cudaStreamWaitEvent(st, ctx->evTrainTrace, 0));
    }
    float lossScale = P.widthFactor * P.heightFactor;   // ONE_OVER_PIXEL_COUNT (nrc-train.comp:210)
    if (ctx->precision == NRC_MLP_BF16_TC)
        NRC_TRY(launch_train_tc(r->mlp, r->mrhe, r->tp.pos, r->tp.dir, r->tp.target, n, lossScale, nullptr, st, deferred));
    else
        NRC_TRY(launch_train_fp32(r->mlp, r->mrhe, r->tp.pos, r->tp.dir, r->tp.target, n, lossScale, nullptr, st, deferred));
    if (ahead) {
        NRC_CUDA(cudaEventRecord(ctx->evTrainFree[b], st));
        r->trainPending[b] = true;
    }
    return NRC_OK;
}

static nrc_status drain_streams(nrc_ctx* ctx)
{
    NRC_CUDA(cudaStreamSynchronize(ctx->stream5));
    for (int i = 0; i < NRC_TRAIN_WALK_STREAMS; i++) NRC_CUDA(cudaStreamSynchronize(ctx->streamWalk[i]));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream4));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream5));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream2));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(ctx->streamCopy));
    return NRC_OK;
}

// Enqueue one render pass with the parameters in r->lastP into the other buffer set.
static nrc_status render_pass(nrc_renderer* r)
{
    nrc_ctx* ctx = r->ctx;
    const TraceParams& P = r->lastP;
    const uint32_t rows = part_count(r->y0, r->y1, r->ystride);
    const uint32_t spp = r->lastSpp;
    static int noOverlap = -1;
    if (noOverlap < 0) { const char* e = getenv("NRC_NO_OVERLAP"); noOverlap = (e && atoi(e)) ? 1 : 0; }
    cudaStream_t s2 = noOverlap ? ctx->stream : ctx->stream2;
    const int b = noOverlap ? r->renderBuf : ((r->renderBuf + 1) % NRC_RENDER_SETS);
    use_render_set(r, b);
    uint32_t* sc = r->scalars + 8 * b;
    unsigned long long* tapc = r->taps + 3 * b;
    // the composite that last read this set (two frames ago) must be done before the walk overwrites it
    if (s2 != ctx->stream && r->setPending[b]) NRC_CUDA(cudaStreamWaitEvent(s2, r->evSetFree[b], 0));
    NRC_CUDA(cudaMemsetAsync(sc + 0, 0, 2 * sizeof(uint32_t), s2));
    NRC_CUDA(cudaMemsetAsync(sc + 4, 0, 2 * sizeof(uint32_t), s2));   // query count, spilled paths of the render walk
    NRC_CUDA(cudaMemsetAsync(tapc, 0, sizeof(unsigned long long), s2));
    NRC_TRY(launch_render_walk(ctx, s2, P, r->y0, rows, r->ystride, spp, r->rpool, r->rp, r->background, r->counters));
    // Lights + combine on a stream of their own: the NEXT frame's render walk (stream2, other buffer set) runs beside this frame's
    // lights.  The walk is bound by instruction issue and the lights by L1 gathers, so side by side they finish sooner than one after
    // the other: 2.75 -> 2.65 ms at 1080p, 1.80 -> 1.71 ms at 720p (profiles/r02_carveout.md; NRC_RENDER_LIGHT_STREAM=0: one stream).
    // Round 1 measured no gain from the same split -- with the tracing kernels still inheriting the MLP kernels' shared-memory carve-out.
    static int lightStream = -1;
    if (lightStream < 0) { const char* e = getenv("NRC_RENDER_LIGHT_STREAM"); lightStream = e ? (atoi(e) != 0) : 1; }
    cudaStream_t sl = (lightStream && s2 != ctx->stream) ? ctx->stream5 : s2;
    if (sl != s2) {
        NRC_CUDA(cudaEventRecord(ctx->evRenderWalk, s2));
        NRC_CUDA(cudaStreamWaitEvent(sl, ctx->evRenderWalk, 0));
    }
    NRC_TRY(launch_light(ctx, sl, P, r->rpool, true, tapc, K_LIGHT_RENDER));
    NRC_TRY(launch_render_combine(ctx, sl, P, r->y0, rows, r->ystride, spp, r->rpool, r->rp, r->ql));
    NRC_CUDA(cudaMemcpyAsync(r->hostScalars, sc, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, sl));
    NRC_CUDA(cudaEventRecord(r->evScalars, sl));
    if (sl != ctx->stream) {
        NRC_CUDA(cudaEventRecord(ctx->evJoin, sl));
        NRC_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->evJoin, 0));
    }
    if (r->lastInfer) NRC_TRY(infer_queries(r, ctx->stream));
    // an asynchronous read-back of the previous frame may still be copying out of the frame buffers
    if (r->copyLast >= 0) NRC_CUDA(cudaStreamWaitEvent(ctx->stream, r->evCopyDone[r->copyLast], 0));
    NRC_TRY(launch_composite(ctx, ctx->stream, P, r->y0, rows, r->ystride, spp, r->rp, r->ql, r->background, r->frame, r->frame8, r->queryProbe,
                             sc, tapc, r->totals));
    NRC_CUDA(cudaEventRecord(r->evSetFree[b], ctx->stream));
    r->setPending[b] = true;
    // Only a pool below the worst case can overflow.  The flag is read LATER (resolve_overflow): a host wait here would keep
    // the next frame's train trace from being queued while this frame is still tracing.
    size_t px = (size_t)rows * r->width;   // this rank's pixels: the pool holds only their vertices
    r->overflowPending = (size_t)r->rpool.cap < px * spp * 32;
    return NRC_OK;
}

// The vertex-pool overflow flag of the last frame.  redo: the caller is about to hand that frame's results out, so a truncated
// frame is traced again with the worst-case pool (same parameters; the weights are the frame's own as long as no later
// training was queued, which is the Render() -> read order).  Otherwise the frame was never looked at: just grow the pool.
static nrc_status resolve_overflow(nrc_renderer* r, bool redo)
{
    if (!r->overflowPending) return NRC_OK;
    NRC_CUDA(cudaEventSynchronize(r->evScalars));
    r->overflowPending = false;
    if (r->hostScalars[1] == 0) return NRC_OK;
    size_t cap = (size_t)part_count(r->y0, r->y1, r->ystride) * r->width * r->lastSpp * 32;
    if (cap > 0x7fffffffu) cap = 0x7fffffffu;
    NRC_REQUIRE((size_t)r->rpool.cap < cap, "vertex pool overflow with the worst-case pool");
    NRC_TRY(drain_streams(r->ctx));
    NRC_TRY(alloc_frame(r, r->lastSpp, (uint32_t)cap));
    if (redo) NRC_TRY(render_pass(r));
    return NRC_OK;
}

nrc_status nrc_renderer_resolve_pending(nrc_renderer* r) { return resolve_overflow(r, true); }

extern "C" nrc_status nrc_renderer_get_pool_capacity(nrc_renderer* r, uint64_t* vertices)
{
    NRC_REQUIRE(r && vertices, "null argument");
    *vertices = r->rpool.cap;
    return NRC_OK;
}
static nrc_status resolve_overflow_fwd(nrc_renderer* r) { return resolve_overflow(r, false); }

// nrc-forward.{vert,frag} over this rank's rows.  The trace part runs on stream2 and may overlap whatever is
// still queued on the main stream (training, steps); inference waits for both.
extern "C" nrc_status nrc_frame_render(nrc_renderer* r)
{
    NRC_REQUIRE(r, "null renderer");
    nrc_ctx* ctx = r->ctx;
    const uint32_t rows = part_count(r->y0, r->y1, r->ystride);
    if (rows == 0) return NRC_OK;
    NRC_TRY(resolve_overflow(r, false));   // the previous frame (costs nothing: this frame's trace queues behind that one anyway)
    const nrc_volume_params& v = r->vol->params;
    const bool useNN = (v.useNN == 1u);
    uint32_t spp = (uint32_t)(useNN ? v.withNnSpp : v.noNnSpp);
    size_t px = (size_t)rows * r->width;   // the vertex pool is sized for this rank's rows
    size_t wantCap = px * spp * (useNN ? 4 : 32);
    if (useNN && r->poolCapHint) wantCap = r->poolCapHint < px * spp * 32 ? r->poolCapHint : px * spp * 32;
    if (useNN && !r->poolCapHint) {
        // No capacity set by the caller: the worst-case pool (32 lit vertices per path, 128 B each, three sets: 25.5 GB at 1080p) when that is a
        // small part of this GPU's memory -- it cannot overflow, so no frame ever waits on the host for its overflow flag (e2e 2.75
        // instead of 2.83 ms at 1080p, 1.77 instead of 2.02 ms at 720p) -- else 4 vertices per path with the deferred check.
        if (r->autoPaths != px * spp) {
            size_t freeB = 0, totalB = 0;
            NRC_CUDA(cudaMemGetInfo(&freeB, &totalB));
            const size_t need = px * spp * 32 * 128 * NRC_RENDER_SETS;
            const size_t have = (size_t)r->rpool.cap * 128 * NRC_RENDER_SETS;   // what the current pools give back when they are replaced
            r->autoWorst = need <= totalB / 6 && need <= (freeB + have) / 2;
            r->autoPaths = px * spp;
        }
        if (r->autoWorst) wantCap = px * spp * 32;
    }
    if (wantCap > 0x7fffffffu) wantCap = 0x7fffffffu;
    if (spp != r->sppAlloc || r->rpool.cap < wantCap) {
        NRC_TRY(drain_streams(ctx));
        NRC_TRY(alloc_frame(r, spp, (uint32_t)wantCap));
    }
    r->lastP = make_params(r);
    r->lastSpp = spp;
    r->lastInfer = useNN && v.showNonNN == 0u;
    return render_pass(r);
}

extern "C" nrc_status nrc_frame_train(nrc_renderer* r) { return frame_train(r, nullptr); }

// Render(): train -> nrc-step -> mrhe-step -> render, like the recorded command buffer.  The sum of the train kernel's per-CTA partial
// gradients and both optimiser steps are ONE launch where that is exactly what the separate kernels would do (steps.cu).
extern "C" nrc_status nrc_render(nrc_renderer* r)
{
    NRC_REQUIRE(r, "null renderer");
    int deferred[2] = {-1, 0};
    NRC_TRY(frame_train(r, deferred));
    if (deferred[0] < 0) {                         // nothing to train on this rank
        NRC_TRY(nrc_optimizer_step(r->mlp, r->mrhe));
    } else if (reduce_step_fusable(r->mlp, r->mrhe, deferred[1] != 0)) {
        NRC_TRY(launch_reduce_step(r->mlp, r->mrhe, deferred[0], r->ctx->stream));
    } else {
        NRC_TRY(launch_reduce_partials(r->mlp, r->mrhe, deferred[0], deferred[1] != 0, r->ctx->stream));
        NRC_TRY(nrc_optimizer_step(r->mlp, r->mrhe));
    }
    NRC_TRY(nrc_frame_render(r));
    return NRC_OK;
}

// ---- read-back ---------------------------------------------------------------------------------------------------------------
extern "C" size_t nrc_frame_bytes(nrc_renderer* r, nrc_frame_format fmt)
{
    if (!r) return 0;
    size_t px = (size_t)r->width * r->height;
    return fmt == NRC_FRAME_RGBA32F ? px * 16 : px * 4;
}

extern "C" void* nrc_frame_ptr(nrc_renderer* r, nrc_frame_format fmt)
{
    if (!r) return nullptr;
    return fmt == NRC_FRAME_RGBA32F ? (void*)r->frame : (void*)r->frame8;
}

static nrc_status read_back(nrc_ctx* ctx, void* dst, const void* src, size_t bytes)
{
    cudaMemcpyKind k = nrc_is_device_ptr(dst) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    NRC_CUDA(cudaMemcpyAsync(dst, src, bytes, k, ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream));
    return NRC_OK;
}

extern "C" nrc_status nrc_read_frame(nrc_renderer* r, void* dst, nrc_frame_format fmt)
{
    NRC_REQUIRE(r && dst, "null argument");
    NRC_REQUIRE(fmt == NRC_FRAME_RGBA32F || fmt == NRC_FRAME_BGRA8, "unknown frame format");
    NRC_TRY(resolve_overflow(r, true));
    return read_back(r->ctx, dst, nrc_frame_ptr(r, fmt), nrc_frame_bytes(r, fmt));
}

// This rank's rows only, compacted: dst[k][width] = frame row y0 + k*ystride.
extern "C" nrc_status nrc_read_frame_partition(nrc_renderer* r, void* dst, nrc_frame_format fmt)
{
    NRC_REQUIRE(r && dst, "null argument");
    NRC_REQUIRE(fmt == NRC_FRAME_RGBA32F || fmt == NRC_FRAME_BGRA8, "unknown frame format");
    uint32_t rows = part_count(r->y0, r->y1, r->ystride);
    if (rows == 0) return NRC_OK;
    NRC_TRY(resolve_overflow(r, true));
    size_t rowBytes = (size_t)r->width * (fmt == NRC_FRAME_RGBA32F ? 16 : 4);
    const char* src = (const char*)nrc_frame_ptr(r, fmt) + (size_t)r->y0 * rowBytes;
    cudaMemcpyKind k = nrc_is_device_ptr(dst) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    NRC_CUDA(cudaMemcpy2DAsync(dst, rowBytes, src, rowBytes * r->ystride, rowBytes, rows, k, r->ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(r->ctx->stream));
    return NRC_OK;
}

// The same copy, queued on the copy stream behind everything the main stream holds now; returns at once.  dst must be
// page-locked host memory (or device memory) that stays valid until nrc_readback_wait(ticket) returns.  The next frame's
// composite waits for the copy, so the frame buffers are never overwritten under it.
extern "C" nrc_status nrc_read_frame_partition_async(nrc_renderer* r, void* dst, nrc_frame_format fmt, int* ticket)
{
    NRC_REQUIRE(r && dst && ticket, "null argument");
    NRC_REQUIRE(fmt == NRC_FRAME_RGBA32F || fmt == NRC_FRAME_BGRA8, "unknown frame format");
    nrc_ctx* ctx = r->ctx;
    NRC_TRY(resolve_overflow(r, true));   // waits for this frame's TRACE only (long done in steady state), not for the frame
    int slot = r->copyNext;
    r->copyNext = (r->copyNext + 1) & 3;
    uint32_t rows = part_count(r->y0, r->y1, r->ystride);
    NRC_CUDA(cudaEventRecord(r->evCopyReady, ctx->stream));
    NRC_CUDA(cudaStreamWaitEvent(ctx->streamCopy, r->evCopyReady, 0));
    if (rows) {
        size_t rowBytes = (size_t)r->width * (fmt == NRC_FRAME_RGBA32F ? 16 : 4);
        const char* src = (const char*)nrc_frame_ptr(r, fmt) + (size_t)r->y0 * rowBytes;
        cudaMemcpyKind k = nrc_is_device_ptr(dst) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
        NRC_CUDA(cudaMemcpy2DAsync(dst, rowBytes, src, rowBytes * r->ystride, rowBytes, rows, k, ctx->streamCopy));
    }
    NRC_CUDA(cudaEventRecord(r->evCopyDone[slot], ctx->streamCopy));
    r->copyLast = slot;
    *ticket = slot;
    return NRC_OK;
}

extern "C" nrc_status nrc_readback_wait(nrc_renderer* r, int ticket)
{
    NRC_REQUIRE(r && ticket >= 0 && ticket < 4, "bad ticket");
    NRC_CUDA(cudaEventSynchronize(r->evCopyDone[ticket]));
    return nrc_ctx_check_dp(r->ctx);
}

extern "C" size_t nrc_frame_partition_bytes(nrc_renderer* r, nrc_frame_format fmt)
{
    if (!r) return 0;
    return (size_t)part_count(r->y0, r->y1, r->ystride) * r->width * (fmt == NRC_FRAME_RGBA32F ? 16 : 4);
}

extern "C" nrc_status nrc_read_render_probes(nrc_renderer* r, uint32_t* counters, float* query)
{
    NRC_REQUIRE(r, "null renderer");
    NRC_TRY(resolve_overflow(r, true));
    size_t px = (size_t)r->width * r->height;
    if (counters) NRC_TRY(read_back(r->ctx, counters, r->counters, px * 4 * sizeof(uint32_t)));
    if (query) NRC_TRY(read_back(r->ctx, query, r->queryProbe, px * 8 * sizeof(float)));
    return NRC_OK;
}

__global__ void k_pack_records(const float* __restrict__ pos, const float* __restrict__ dir, const float* __restrict__ target, uint32_t n,
                               float* __restrict__ rec)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    for (int k = 0; k < 3; k++) {
        rec[9 * (size_t)i + k] = pos[3 * (size_t)i + k];
        rec[9 * (size_t)i + 3 + k] = dir[3 * (size_t)i + k];
        rec[9 * (size_t)i + 6 + k] = target[3 * (size_t)i + k];
    }
}

// records / trainCounters hold trainW*trainH rows; this rank's rows are [i0, i1), the rest are left untouched.
extern "C" nrc_status nrc_read_train_probes(nrc_renderer* r, float* records, uint32_t* trainCounters)
{
    NRC_REQUIRE(r, "null renderer");
    uint32_t n = part_count(r->i0, r->i1, r->istride);
    if (n == 0) return NRC_OK;
    nrc_ctx* ctx = r->ctx;
    if (r->istride > 1) {   // strided shard: scatter rows i0 + k*istride
        cudaMemcpyKind k = nrc_is_device_ptr(records ? (void*)records : (void*)trainCounters) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
        if (records) {
            void* tmp;
            NRC_TRY(nrc_scratch(ctx, 5, (size_t)n * 9 * sizeof(float), &tmp));
            NRC_LAUNCH(ctx, ctx->stream, K_PACK_RECORDS,
                       k_pack_records<<<(n + 255) / 256, 256, 0, ctx->stream>>>(r->tp.pos, r->tp.dir, r->tp.target, n, (float*)tmp));
            NRC_CUDA(cudaMemcpy2DAsync(records + (size_t)r->i0 * 9, 36 * (size_t)r->istride, tmp, 36, 36, n, k, ctx->stream));
        }
        if (trainCounters)
            NRC_CUDA(cudaMemcpy2DAsync(trainCounters + (size_t)r->i0 * 4, 16 * (size_t)r->istride, r->trainCounters, 16, 16, n, k, ctx->stream));
        NRC_CUDA(cudaStreamSynchronize(ctx->stream));
        return NRC_OK;
    }
    if (records) {
        void* tmp;
        NRC_TRY(nrc_scratch(ctx, 5, (size_t)n * 9 * sizeof(float), &tmp));
        NRC_LAUNCH(ctx, ctx->stream, K_PACK_RECORDS,
                   k_pack_records<<<(n + 255) / 256, 256, 0, ctx->stream>>>(r->tp.pos, r->tp.dir, r->tp.target, n, (float*)tmp));
        NRC_TRY(read_back(ctx, records + (size_t)r->i0 * 9, tmp, (size_t)n * 9 * sizeof(float)));
    }
    if (trainCounters) NRC_TRY(read_back(ctx, trainCounters + (size_t)r->i0 * 4, r->trainCounters, (size_t)n * 4 * sizeof(uint32_t)));
    return NRC_OK;
}

extern "C" nrc_status nrc_frame_totals(nrc_renderer* r, uint64_t totals[6], int reset)
{
    NRC_REQUIRE(r && totals, "null argument");
    nrc_ctx* ctx = r->ctx;
    for (int i = 0; i < NRC_TRAIN_WALK_STREAMS; i++) NRC_CUDA(cudaStreamSynchronize(ctx->streamWalk[i]));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream4));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream5));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream2));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream));
    unsigned long long t[6];
    NRC_CUDA(cudaMemcpy(t, r->totals, sizeof(t), cudaMemcpyDeviceToHost));
    for (int i = 0; i < 6; i++) totals[i] = t[i];
    if (reset) NRC_CUDA(cudaMemset(r->totals, 0, sizeof(t)));
    return NRC_OK;
}

extern "C" nrc_status nrc_frame_stats(nrc_renderer* r, uint64_t* inferQueries, uint64_t* trainSamples, uint64_t* densityTaps)
{
    NRC_REQUIRE(r, "null renderer");
    nrc_ctx* ctx = r->ctx;
    NRC_TRY(resolve_overflow(r, true));
    for (int i = 0; i < NRC_TRAIN_WALK_STREAMS; i++) NRC_CUDA(cudaStreamSynchronize(ctx->streamWalk[i]));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream4));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream5));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream2));
    NRC_CUDA(cudaMemcpyAsync(r->hostScalars, r->scalars + 8 * r->renderBuf, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    NRC_CUDA(cudaMemcpyAsync(r->hostTaps, r->taps, 16 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    NRC_CUDA(cudaStreamSynchronize(ctx->stream));
    if (inferQueries) *inferQueries = r->hostScalars[4];
    if (trainSamples) *trainSamples = part_count(r->i0, r->i1, r->istride);
    // + one tap per walk step is not counted here: densityTaps are the light kernels' taps (the bulk)
    if (densityTaps) *densityTaps = (uint64_t)r->hostTaps[3 * r->renderBuf] + (uint64_t)r->hostTaps[12 + r->trainBuf];
    return NRC_OK;
}
