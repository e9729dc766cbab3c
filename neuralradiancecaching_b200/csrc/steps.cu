// steps.cu -- optimiser steps and parameter initialisation.
//
// nrc-step.comp:200-406  : g' = (bad(g) ? 1000*sign(g) : g) - wd*w ; m = (1-b1)*g' + b1*m ; w += lr*m ;
//                          bad(w) -> 0 for weights only (A14) ; delta = 0.      24 B/param, 20 995 params.
// mrhe-step.comp:27-74   : g' = (bad(g) ? 100*sign(g) : g) - wd*w ; w += lr*g' ; bad(w) -> 0 ; delta = 0.
//                          16 B/param streamed with 128-bit accesses; HBM-bound at T = 2^22 (2 GiB / step).
// Both are pure streaming kernels: grid sized in multiples of the SM count, float4 accesses, no reuse.
#include <cuda_bf16.h>
#include <stdlib.h>
#include "nrc_internal.cuh"
#include "tc_ptx.cuh"

__device__ __forceinline__ bool bad_value(float x) { return (x != x) || (fabsf(x) == INFINITY) || (fabsf(x) > 1000.0f); }
__device__ __forceinline__ float sign_nan0(float x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : 0.0f); }

// Also refreshes the transposed weight image Wt[l][in][out] used by the forward GEMMs and clears mseLoss? No:
// mseLoss (grad[20995]) is reset by ResetStats, as in the reference (src/main.cpp:215).
__global__ void __launch_bounds__(256) k_nrc_step(float* __restrict__ params, float* __restrict__ grad, float* __restrict__ momentum,
                                                  float* __restrict__ paramsT, unsigned short* __restrict__ imgBf16, float lr, float wd,
                                                  float beta1)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= NRC_MLP_PARAM_COUNT) return;
    float w = params[i];
    float g = grad[i];
    if (bad_value(g)) g = __fmul_rn(1000.0f, sign_nan0(g));
    g = __fadd_rn(g, __fmul_rn(-w, wd));
    float d = __fadd_rn(__fmul_rn(__fsub_rn(1.0f, beta1), g), __fmul_rn(beta1, momentum[i]));
    momentum[i] = d;
    float nw = __fadd_rn(w, __fmul_rn(d, lr));
    if (i < NRC_MLP_WEIGHT_COUNT && bad_value(nw)) nw = 0.0f;
    params[i] = nw;
    grad[i] = 0.0f;
    if (i < 5 * 4096) {
        int l = i >> 12, rem = i & 4095, o = rem >> 6, k = rem & 63;
        paramsT[l * 4096 + k * 64 + o] = nw;
    }
    if (imgBf16 && i < NRC_MLP_WEIGHT_COUNT) tc_store_weight(imgBf16, i, nw);   // operand images of the tcgen05 path (tc_ptx.cuh)
}

// Opt-in Adam (the oracle's NrcStep with S.adam): the same scrub / weight decay / first moment as above, then
// v = b2*v + (1-b2)*g'^2 and w += lr * (m*c1) / (sqrt(v*c2) + eps) with the bias corrections c = 1/(1 - beta^t) computed on the host.
__global__ void __launch_bounds__(256) k_nrc_step_adam(float* __restrict__ params, float* __restrict__ grad, float* __restrict__ momentum,
                                                       float* __restrict__ second, float* __restrict__ paramsT, unsigned short* __restrict__ imgBf16,
                                                       float lr, float wd, float beta1, float beta2, float eps, float c1, float c2)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= NRC_MLP_PARAM_COUNT) return;
    float w = params[i];
    float g = grad[i];
    if (bad_value(g)) g = __fmul_rn(1000.0f, sign_nan0(g));
    g = __fadd_rn(g, __fmul_rn(-w, wd));
    float d = __fadd_rn(__fmul_rn(__fsub_rn(1.0f, beta1), g), __fmul_rn(beta1, momentum[i]));
    momentum[i] = d;
    float v = __fadd_rn(__fmul_rn(beta2, second[i]), __fmul_rn(__fsub_rn(1.0f, beta2), __fmul_rn(g, g)));
    second[i] = v;
    d = __fdiv_rn(__fmul_rn(d, c1), __fadd_rn(__fsqrt_rn(__fmul_rn(v, c2)), eps));
    float nw = __fadd_rn(w, __fmul_rn(d, lr));
    if (i < NRC_MLP_WEIGHT_COUNT && bad_value(nw)) nw = 0.0f;
    params[i] = nw;
    grad[i] = 0.0f;
    if (i < 5 * 4096) {
        int l = i >> 12, rem = i & 4095, o = rem >> 6, k = rem & 63;
        paramsT[l * 4096 + k * 64 + o] = nw;
    }
    if (imgBf16 && i < NRC_MLP_WEIGHT_COUNT) tc_store_weight(imgBf16, i, nw);
}

__device__ __forceinline__ float mrhe_update_adam(float w, float g, float& m1, float& m2, float lr, float wd, float beta1, float beta2, float eps,
                                                  float c1, float c2)
{
    if (bad_value(g)) g = __fmul_rn(100.0f, sign_nan0(g));
    g = __fadd_rn(g, __fmul_rn(-w, wd));
    m1 = __fadd_rn(__fmul_rn(beta1, m1), __fmul_rn(__fsub_rn(1.0f, beta1), g));
    m2 = __fadd_rn(__fmul_rn(beta2, m2), __fmul_rn(__fsub_rn(1.0f, beta2), __fmul_rn(g, g)));
    float d = __fdiv_rn(__fmul_rn(m1, c1), __fadd_rn(__fsqrt_rn(__fmul_rn(m2, c2)), eps));
    float nw = __fadd_rn(w, __fmul_rn(d, lr));
    return bad_value(nw) ? 0.0f : nw;
}

// 32 B/param streamed (w, delta, m1, m2 read; w, delta, m1, m2 written)
__global__ void __launch_bounds__(256) k_mrhe_step_adam(float4* __restrict__ tables, float4* __restrict__ delta, float4* __restrict__ m1,
                                                        float4* __restrict__ m2, size_t n4, float lr, float wd, float beta1, float beta2, float eps,
                                                        float c1, float c2)
{
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 w = tables[i], g = delta[i], a = m1[i], b = m2[i];
        w.x = mrhe_update_adam(w.x, g.x, a.x, b.x, lr, wd, beta1, beta2, eps, c1, c2);
        w.y = mrhe_update_adam(w.y, g.y, a.y, b.y, lr, wd, beta1, beta2, eps, c1, c2);
        w.z = mrhe_update_adam(w.z, g.z, a.z, b.z, lr, wd, beta1, beta2, eps, c1, c2);
        w.w = mrhe_update_adam(w.w, g.w, a.w, b.w, lr, wd, beta1, beta2, eps, c1, c2);
        tables[i] = w; m1[i] = a; m2[i] = b;
        delta[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
}

__device__ __forceinline__ float mrhe_update(float w, float g, float lr, float wd)
{
    if (bad_value(g)) g = __fmul_rn(100.0f, sign_nan0(g));
    g = __fadd_rn(g, __fmul_rn(-w, wd));
    float nw = __fadd_rn(w, __fmul_rn(g, lr));
    return bad_value(nw) ? 0.0f : nw;
}

__global__ void __launch_bounds__(256) k_mrhe_step(float4* __restrict__ tables, float4* __restrict__ delta, size_t n4, float lr, float wd)
{
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 w = tables[i];
        float4 g = delta[i];
        w.x = mrhe_update(w.x, g.x, lr, wd);
        w.y = mrhe_update(w.y, g.y, lr, wd);
        w.z = mrhe_update(w.z, g.z, lr, wd);
        w.w = mrhe_update(w.w, g.w, lr, wd);
        tables[i] = w;
        delta[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
}

// ---- one launch for {k_reduce_partials, k_nrc_step, k_mrhe_step} of a single-GPU frame ---------------------------------------------------
// Thread i < 20 995: the fixed-order sum of the train kernel's per-CTA partial gradients, added to what the arena already holds, then
// nrc-step.comp on it (the arithmetic of k_nrc_step, operation for operation).  Thread 20 995: the loss accumulator.  Threads from
// NRC_HIST_FIRST on: entry h of the delta table's head = the fixed-order sum of the CTAs' histograms (both features, A9), then
// mrhe-step.comp on it.  Only valid where the separate kernels would do exactly this: momentum-SGD steps (no Adam), the shared-memory
// histogram scatter (so the head is covered entry for entry), and a table tail that is a fixed point of the step (nrc_mrhe::tailClean).
#define NRC_HIST_FIRST 21248   // as in mlp_fp32.cu: first histogram thread, a multiple of the block size past the arena
__global__ void __launch_bounds__(256) k_reduce_step(const float* __restrict__ partial, int nb, float* __restrict__ grad, float* __restrict__ params,
                                                     float* __restrict__ momentum, float* __restrict__ paramsT, unsigned short* __restrict__ imgBf16,
                                                     float lr, float wd, float beta1, const float* __restrict__ hist, uint32_t T,
                                                     float2* __restrict__ delta, float2* __restrict__ tables, float mlr, float mwd)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i <= NRC_MLP_PARAM_COUNT) {
        float s = 0.0f;
        for (int b = 0; b < nb; b++) s += partial[(size_t)b * NRC_PARTIAL_STRIDE + i];
        if (i == NRC_MLP_PARAM_COUNT) { grad[i] += s; return; }   // mseLoss
        const float w = params[i];
        float g = grad[i] + s;
        if (bad_value(g)) g = __fmul_rn(1000.0f, sign_nan0(g));
        g = __fadd_rn(g, __fmul_rn(-w, wd));
        const float d = __fadd_rn(__fmul_rn(__fsub_rn(1.0f, beta1), g), __fmul_rn(beta1, momentum[i]));
        momentum[i] = d;
        float nw = __fadd_rn(w, __fmul_rn(d, lr));
        if (i < NRC_MLP_WEIGHT_COUNT && bad_value(nw)) nw = 0.0f;
        params[i] = nw;
        grad[i] = 0.0f;
        if (i < 5 * 4096) {
            int l = i >> 12, rem = i & 4095, o = rem >> 6, k = rem & 63;
            paramsT[l * 4096 + k * 64 + o] = nw;
        }
        if (imgBf16 && i < NRC_MLP_WEIGHT_COUNT) tc_store_weight(imgBf16, i, nw);
        return;
    }
    const uint32_t h = (uint32_t)i - NRC_HIST_FIRST;
    if (i < NRC_HIST_FIRST || h >= T) return;
    float s = 0.0f;
    for (int b = 0; b < nb; b++) s += hist[(size_t)b * T + h];
    float2 d = delta[h];
    d.x += s; d.y += s;
    float2 w = tables[h];
    w.x = mrhe_update(w.x, d.x, mlr, mwd);
    w.y = mrhe_update(w.y, d.y, mlr, mwd);
    tables[h] = w;
    delta[h] = make_float2(0.0f, 0.0f);
}

// true if the fused launch is exactly {reduce, nrc-step, mrhe-step} for the state the objects are in (see above)
bool reduce_step_fusable(const nrc_mlp* mlp, const nrc_mrhe* m, bool hist)
{
    // Measured on the 1080p bench frame (alternating runs, same box): 2.938 / 2.933 ms fused against 2.906 / 2.875 ms with the three
    // separate launches -- no gain (the frame is throughput-bound; two launch gaps on the main stream are not on its critical path).
    // So it is opt-in: NRC_FUSED_STEP=1.
    const char* e = getenv("NRC_FUSED_STEP");   // read at every frame (tests switch it)
    const bool on = e && atoi(e);
    return on && hist && !mlp->adam && !m->adam && !m->fixed && m->tailClean && m->wd == 0.0f && (size_t)2 * m->T < m->count;
}

nrc_status launch_reduce_step(nrc_mlp* mlp, nrc_mrhe* m, int blocks, cudaStream_t st)
{
    const unsigned threads = NRC_HIST_FIRST + m->T;
    NRC_LAUNCH(mlp->ctx, st, K_REDUCE_STEP,
               k_reduce_step<<<(threads + 255) / 256, 256, 0, st>>>(mlp->partial, blocks, mlp->grad, mlp->params, mlp->momentum, mlp->paramsT,
                                                                    mlp->paramsBf16, mlp->lr, mlp->wd, mlp->beta1, mlp->histPartial, m->T,
                                                                    reinterpret_cast<float2*>(m->delta), reinterpret_cast<float2*>(m->tables), m->lr, m->wd));
    return NRC_OK;
}

// Counter-based N(0,1)*scale: splitmix64 per element + Box-Muller.  (The reference seeds from
// std::random_device, src/NeuralRadianceCache.cpp:629 / src/MRHE.cpp:127: not reproducible; parity runs set
// parameters explicitly.)
__device__ __forceinline__ uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

__global__ void __launch_bounds__(256) k_init_normal(float* __restrict__ dst, size_t n, float scale, uint64_t seed)
{
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        uint64_t h = splitmix64(seed * 0xD1342543DE82EF95ull + i);
        float u1 = ((float)((h >> 40) + 1)) * (1.0f / 16777217.0f);
        float u2 = ((float)((h >> 8) & 0xffffffu)) * (1.0f / 16777216.0f);
        float r = sqrtf(-2.0f * logf(u1));
        dst[i] = r * cosf(6.283185307179586f * u2) * scale;
    }
}

nrc_status launch_nrc_step(nrc_mlp* mlp, cudaStream_t st)
{
    if (mlp->adam) {
        mlp->steps++;
        const float c1 = 1.0f / (1.0f - powf(mlp->beta1, (float)mlp->steps)), c2 = 1.0f / (1.0f - powf(mlp->beta2, (float)mlp->steps));
        NRC_LAUNCH(mlp->ctx, st, K_NRC_STEP,
                   k_nrc_step_adam<<<(NRC_MLP_PARAM_COUNT + 255) / 256, 256, 0, st>>>(mlp->params, mlp->grad, mlp->momentum, mlp->second, mlp->paramsT,
                                                                                     mlp->paramsBf16, mlp->lr, mlp->wd, mlp->beta1, mlp->beta2,
                                                                                     mlp->eps, c1, c2));
        return NRC_OK;
    }
    NRC_LAUNCH(mlp->ctx, st, K_NRC_STEP,
               k_nrc_step<<<(NRC_MLP_PARAM_COUNT + 255) / 256, 256, 0, st>>>(mlp->params, mlp->grad, mlp->momentum, mlp->paramsT,
                                                                            mlp->paramsBf16, mlp->lr, mlp->wd, mlp->beta1));
    return NRC_OK;
}

// entries beyond which the step may skip (see nrc_mrhe::tailClean): the whole table unless the tail is a proven fixed point
static size_t mrhe_step_extent(const nrc_mrhe* m)
{
    const size_t head = nrc_mrhe_grad_extent(const_cast<nrc_mrhe*>(m));
    static int noSkip = -1;   // NRC_MRHE_STEP_FULL=1: always stream the whole table like the reference's dispatch (measurements)
    if (noSkip < 0) { const char* e = getenv("NRC_MRHE_STEP_FULL"); noSkip = (e && atoi(e)) ? 1 : 0; }
    if (!noSkip && m->tailClean && m->wd == 0.0f && !m->adam && !m->fixed && head < m->count) return head;
    return m->count;
}

nrc_status launch_mrhe_step(nrc_mrhe* m, cudaStream_t st)
{
    const size_t extent = mrhe_step_extent(m);
    size_t n4 = extent / 4;   // L*T*F is a multiple of 4 for every supported shape (F == 2, T even)
    size_t want = (n4 + 255) / 256;
    size_t cap = (size_t)m->ctx->smCount * 8;
    unsigned blocks = (unsigned)(want < cap ? want : cap);
    if (blocks == 0) return NRC_OK;
    if (m->adam) {
        m->steps++;
        const float c1 = 1.0f / (1.0f - powf(m->beta1, (float)m->steps)), c2 = 1.0f / (1.0f - powf(m->beta2, (float)m->steps));
        NRC_LAUNCH(m->ctx, st, K_MRHE_STEP,
                   k_mrhe_step_adam<<<blocks, 256, 0, st>>>(reinterpret_cast<float4*>(m->tables), reinterpret_cast<float4*>(m->delta),
                                                            reinterpret_cast<float4*>(m->m1), reinterpret_cast<float4*>(m->m2), n4, m->lr, m->wd,
                                                            m->beta1, m->beta2, m->eps, c1, c2));
        return NRC_OK;
    }
    NRC_LAUNCH(m->ctx, st, K_MRHE_STEP,
               k_mrhe_step<<<blocks, 256, 0, st>>>(reinterpret_cast<float4*>(m->tables), reinterpret_cast<float4*>(m->delta), n4, m->lr, m->wd));
    if (extent == m->count && m->wd == 0.0f && !m->fixed) m->tailClean = true;   // every entry scrubbed, every delta zero: the tail is a fixed point now
    return NRC_OK;
}

// tail [firstFloat, count) of the tables: the part of mrhe-step the fused data-parallel kernel (dp.cu) leaves to this one
nrc_status launch_mrhe_step_range(nrc_mrhe* m, size_t firstFloat, cudaStream_t st)
{
    if (firstFloat >= m->count) return NRC_OK;
    if (firstFloat >= mrhe_step_extent(m)) return NRC_OK;   // the tail is a fixed point of the step (nrc_mrhe::tailClean)
    size_t n4 = (m->count - firstFloat) / 4;
    size_t want = (n4 + 255) / 256;
    size_t cap = (size_t)m->ctx->smCount * 8;
    unsigned blocks = (unsigned)(want < cap ? want : cap);
    if (blocks == 0) return NRC_OK;
    if (m->adam) {   // the fused kernel has already counted this step (dp.cu): same bias corrections, no second increment
        const float c1 = 1.0f / (1.0f - powf(m->beta1, (float)m->steps)), c2 = 1.0f / (1.0f - powf(m->beta2, (float)m->steps));
        NRC_LAUNCH(m->ctx, st, K_MRHE_STEP,
                   k_mrhe_step_adam<<<blocks, 256, 0, st>>>(reinterpret_cast<float4*>(m->tables + firstFloat), reinterpret_cast<float4*>(m->delta + firstFloat),
                                                            reinterpret_cast<float4*>(m->m1 + firstFloat), reinterpret_cast<float4*>(m->m2 + firstFloat), n4,
                                                            m->lr, m->wd, m->beta1, m->beta2, m->eps, c1, c2));
        return NRC_OK;
    }
    NRC_LAUNCH(m->ctx, st, K_MRHE_STEP,
               k_mrhe_step<<<blocks, 256, 0, st>>>(reinterpret_cast<float4*>(m->tables + firstFloat),
                                                   reinterpret_cast<float4*>(m->delta + firstFloat), n4, m->lr, m->wd));
    if (firstFloat <= nrc_mrhe_grad_extent(m) && m->wd == 0.0f && !m->fixed && !m->adam) m->tailClean = true;
    return NRC_OK;
}

nrc_status launch_init_normal(nrc_ctx* ctx, float* dst, size_t n, float scale, uint64_t seed)
{
    if (n == 0) return NRC_OK;
    size_t want = (n + 255) / 256;
    size_t cap = (size_t)ctx->smCount * 8;
    unsigned blocks = (unsigned)(want < cap ? want : cap);
    NRC_LAUNCH(ctx, ctx->stream, K_INIT_NORMAL, k_init_normal<<<blocks, 256, 0, ctx->stream>>>(dst, n, scale, seed));
    return NRC_OK;
}
