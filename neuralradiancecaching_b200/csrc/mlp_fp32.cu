// mlp_fp32.cu -- fused MRHE/OneBlob encode + 64-wide MLP, fp32 on the CUDA cores.
//
// This is the 1e-3 parity path of SURVEY.md 8a rows a11 (Forward, nrc-forward.frag:503-534), a12 (Backprop,
// nrc-train.comp:743-945, 985-1028) and a13 (BackpropMrhe, :947-983).  The reference runs one thread per
// sample with 7 x 64 private floats and one global float atomic per weight per sample; here a CTA owns a
// tile of 64 samples and every layer is a 64x64x64 shared-memory GEMM:
//
//   C[m][n] = sum_k P[k][m] * Q[k][n]      (both operands "k-major": the reduction index is the slow one)
//
//   forward   : act_{l+1}[r][c] = sum_k  Xt[k][r]   * Wt_l[k][c]     (k = input feature)
//   back-data : err_l[r][c]     = sum_k  Et[k][r]   * W_{l+1}[k][c]  (k = output feature of layer l+1)
//   back-wgt  : dW_l[o][i]      = sum_r  Es[r][o]   * As_l[r][i]     (k = sample)
//
// so activations/errors are kept feature-major (Xt, Et) for the first two and sample-major (As, Es) for the
// third.  256 threads, 4x4 register tile each, k sequential 0..63 with one fmaf per term: the same summation
// order as the reference's "sum += in * w" loop.  Weight gradients are reduced in-CTA over the 64 samples,
// accumulated into a per-CTA partial arena in L2, and summed over CTAs in fixed order by k_reduce_partials:
// deterministic, and no float atomics on the 20 995 MLP parameters at all.  The hash-table gradient keeps
// the reference's scatter (one red.global.add.v2.f32 per corner).
#include "nrc_internal.cuh"
#include "strict.cuh"
#include "mrhe_scatter.cuh"

#define TILE 64          // samples per CTA tile
#define NRC_HIST_FIRST 21248   // first thread of k_reduce_partials that reduces a histogram entry (a multiple of the block size past the arena)
#define NTHREADS 256

struct MlpDev {
    const float* params;    // packed arena (W row-major [out][in], biases)
    const float* paramsT;   // Wt[l][in][out], l = 0..4
};

// ---- shared device pieces -----------------------------------------------------------------------------

__device__ __forceinline__ void copy16k(float* dst, const float* __restrict__ src)
{
    const float4* s = reinterpret_cast<const float4*>(src);
    float4* d = reinterpret_cast<float4*>(dst);
#pragma unroll
    for (int i = 0; i < 4; i++) d[threadIdx.x + i * NTHREADS] = __ldg(s + threadIdx.x + i * NTHREADS);
}

// acc[i][j] += sum_k P[k*64 + 4*tm + i] * Q[k*64 + 4*tn + j]
__device__ __forceinline__ void gemm64(const float* __restrict__ P, const float* __restrict__ Q, int tm, int tn, float acc[4][4])
{
#pragma unroll 8
    for (int k = 0; k < 64; k++) {
        float4 p = *reinterpret_cast<const float4*>(P + k * 64 + 4 * tm);
        float4 q = *reinterpret_cast<const float4*>(Q + k * 64 + 4 * tn);
        float pv[4] = {p.x, p.y, p.z, p.w};
        float qv[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[i][j] = fmaf(pv[i], qv[j], acc[i][j]);
    }
}

// Encode TILE queries into Xt (feature-major [64][TILE]) and optionally As (sample-major [TILE][64]).
// Thread t: sample r = t % 64, quarter h = t / 64 -> MRHE levels 4h..4h+3 and 8 of the 32 OneBlob features.
// Returns (for h == 0 threads) whether the sample trips the reference's NaN guard on nr0[0..4].
__device__ __forceinline__ void encode_tile(const float* __restrict__ pos, const float* __restrict__ dir, size_t base, size_t n,
                                            const MrheDev& m, float* Xt, float* As)
{
    int r = threadIdx.x & 63, h = threadIdx.x >> 6;
    size_t q = base + r;
    f3 p = mk3(0.0f, 0.0f, 0.0f), d = mk3(0.0f, 0.0f, 1.0f);
    if (q < n) {
        p = mk3(__ldg(pos + 3 * q), __ldg(pos + 3 * q + 1), __ldg(pos + 3 * q + 2));
        d = mk3(__ldg(dir + 3 * q), __ldg(dir + 3 * q + 1), __ldg(dir + 3 * q + 2));
    }
    f3 sky = mk3(m.skySize[0], m.skySize[1], m.skySize[2]);
    f3 np = mrhe_norm_pos(p, sky);
#pragma unroll
    for (int li = 0; li < 4; li++) {
        uint32_t level = 4 * h + li;
        float o0 = 0.0f, o1 = 0.0f;
        if (level < m.L) {
            uint32_t idx[8];
            f3 lerp;
            mrhe_level_indices(p, np, m.res[level], m.T, idx, lerp, m.fixed != 0);
                lerp = mrhe_blend_factors(lerp, m.fixed != 0);
            const float2* tab = reinterpret_cast<const float2*>(m.tables) + (size_t)level * m.T;
            float f0[8], f1[8];
#pragma unroll
            for (int c = 0; c < 8; c++) {
                float2 v = __ldg(tab + idx[c]);
                f0[c] = v.x;
                f1[c] = v.y;
            }
            o0 = mrhe_blend(f0, lerp);
            o1 = mrhe_blend(f1, lerp);
        }
        Xt[(2 * level) * TILE + r] = o0;
        Xt[(2 * level + 1) * TILE + r] = o1;
        if (As) { As[r * 64 + 2 * level] = o0; As[r * 64 + 2 * level + 1] = o1; }
    }
    // OneBlob: quarter h owns bins 8*(h&1)..+7 of theta (h < 2) or phi (h >= 2)
    {
        float ang;
        if (h < 2) {
            ang = NRC_ADD(NRC_DIV(nrc_atan2(d.z, d.x), NRC_PI), 0.5f);
        } else {
            float lenxz = NRC_SQRT(NRC_ADD(NRC_MUL(d.x, d.x), NRC_MUL(d.z, d.z)));
            ang = NRC_ADD(NRC_DIV(nrc_atan2(lenxz, d.y), NRC_PI), 0.5f);
        }
        const float sigma = 0.25f;
        float term1 = NRC_DIV(1.0f, NRC_MUL(sigma, NRC_SQRT(NRC_MUL(2.0f, NRC_PI))));
        int f0 = 32 + ((h >= 2) ? 16 : 0) + 8 * (h & 1);
#pragma unroll
        for (int i = 0; i < 8; i++) {
            float fi = (float)(8 * (h & 1) + i);
            float t2 = NRC_DIV(NRC_SUB(fi, ang), sigma);
            float v = NRC_MUL(term1, nrc_exp(NRC_MUL(NRC_MUL(-0.5f, t2), t2)));
            Xt[(f0 + i) * TILE + r] = v;
            if (As) As[r * 64 + f0 + i] = v;
        }
    }
}

// One hidden layer: Xout = relu(Xin * Wt + b), feature-major in and out; optional sample-major copy.
__device__ __forceinline__ void hidden_layer(const float* Xin, const float* Wt, const float* __restrict__ bias, float* Xout, float* AsOut)
{
    int tm = threadIdx.x & 15, tn = threadIdx.x >> 4;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.0f;
    gemm64(Xin, Wt, tm, tn, acc);
    float b[4];
#pragma unroll
    for (int j = 0; j < 4; j++) b[j] = __ldg(bias + 4 * tn + j);
#pragma unroll
    for (int j = 0; j < 4; j++) {
        float4 col;
        col.x = fmaxf(acc[0][j] + b[j], 0.0f);
        col.y = fmaxf(acc[1][j] + b[j], 0.0f);
        col.z = fmaxf(acc[2][j] + b[j], 0.0f);
        col.w = fmaxf(acc[3][j] + b[j], 0.0f);
        *reinterpret_cast<float4*>(Xout + (4 * tn + j) * TILE + 4 * tm) = col;
        acc[0][j] = col.x; acc[1][j] = col.y; acc[2][j] = col.z; acc[3][j] = col.w;
    }
    if (AsOut) {
#pragma unroll
        for (int i = 0; i < 4; i++)
            *reinterpret_cast<float4*>(AsOut + (4 * tm + i) * 64 + 4 * tn) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
    }
}

// ---- inference -------------------------------------------------------------------------------------------
// smem: W (16 KB) + X0, X1 (16 KB each) = 48 KB -> 4 CTAs/SM.
__global__ void __launch_bounds__(NTHREADS, 4) k_infer_fp32(const float* __restrict__ pos, const float* __restrict__ dir,
                                                            const uint32_t* __restrict__ countDev, size_t nMax, MlpDev w, MrheDev m,
                                                            float* __restrict__ rgb)
{
    extern __shared__ __align__(16) float smem[];
    float* W = smem;
    float* X0 = smem + 4096;
    float* X1 = smem + 8192;
    size_t n = nMax;
    if (countDev) { size_t c = *countDev; n = c < nMax ? c : nMax; }
    size_t tiles = (n + TILE - 1) / TILE;
    for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        size_t base = tile * TILE;
        encode_tile(pos, dir, base, n, m, X0, nullptr);
        float* xin = X0;
        float* xout = X1;
#pragma unroll 1
        for (int l = 0; l < 5; l++) {
            copy16k(W, w.paramsT + l * 4096);
            __syncthreads();
            hidden_layer(xin, W, w.params + NRC_B_OFF(l), xout, nullptr);
            __syncthreads();
            float* t = xin; xin = xout; xout = t;
        }
        // output layer 64 -> 3 (+ ReLU twice, nrc-forward.frag:525-531): thread t < 192: row r = t%64, out j = t/64
        if (threadIdx.x < 192) {
            int r = threadIdx.x & 63, j = threadIdx.x >> 6;
            const float* w5 = w.params + NRC_W_OFF(5) + j * 64;
            float sum = 0.0f;
#pragma unroll 8
            for (int k = 0; k < 64; k++) sum = fmaf(xin[k * TILE + r], __ldg(w5 + k), sum);
            sum = fmaxf(sum + __ldg(w.params + NRC_B_OFF(5) + j), 0.0f);
            if (base + r < n) rgb[3 * (base + r) + j] = sum;
        }
        __syncthreads();
    }
}

// ---- training --------------------------------------------------------------------------------------------
// smem (floats): As[6][64][64] | XE0, XE1 (feature-major) | ES0, ES1 (sample-major) | W | small  = ~178 KB, 1 CTA/SM.
#define TRAIN_SMEM_FLOATS (6 * 4096 + 2 * 4096 + 2 * 4096 + 4096 + 1024)

__device__ __forceinline__ void zero_acc(float acc[4][4])
{
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = 0.0f;
}

__global__ void __launch_bounds__(NTHREADS, 1) k_train_fp32(const float* __restrict__ pos, const float* __restrict__ dir,
                                                            const float* __restrict__ target, const uint32_t* __restrict__ countDev,
                                                            size_t nMax, float scale, MlpDev w, MrheDev m, float* __restrict__ partialAll,
                                                            float* __restrict__ predOut, float* gAll, float* __restrict__ histAll)
{
    extern __shared__ __align__(16) float smem[];
    float* As = smem;                     // [6][64][64]
    float* XE0 = smem + 6 * 4096;
    float* XE1 = XE0 + 4096;
    float* ES0 = XE1 + 4096;
    float* ES1 = ES0 + 4096;
    float* W = ES1 + 4096;
    float* predS = W + 4096;              // [64][4]
    float* e6S = predS + 256;             // [64][4]
    float* lossS = e6S + 256;             // [8]
    int* dropS = reinterpret_cast<int*>(lossS + 8);   // [64]

    const int tid = threadIdx.x;
    const int tm = tid & 15, tn = tid >> 4;
    float* partial = partialAll + (size_t)blockIdx.x * NRC_PARTIAL_STRIDE;
    for (int i = tid; i < NRC_PARTIAL_STRIDE; i += NTHREADS) partial[i] = 0.0f;
    __syncthreads();

    size_t n = nMax;
    if (countDev) { size_t c = *countDev; n = c < nMax ? c : nMax; }
    size_t tiles = (n + TILE - 1) / TILE;
    float lossAcc = 0.0f;   // thread 0 only

    for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const size_t base = tile * TILE;
        // ---- encode (EncodeRay, nrc-train.comp:703-713) ----
        encode_tile(pos, dir, base, n, m, XE0, As);
        __syncthreads();
        // NaN guard on nr0[0..4] (nrc-train.comp:993-999); rows past n are dropped too
        if (tid < 64) {
            const float* a0 = As + tid * 64;
            int bad = (base + tid >= n);
            for (int i = 0; i < 5; i++) bad |= (a0[i] != a0[i]);
            dropS[tid] = bad;
        }
        __syncthreads();
        {
            int r = tid & 63, h = tid >> 6;
            if (dropS[r]) {
                for (int f = 16 * h; f < 16 * h + 16; f++) { XE0[f * TILE + r] = 0.0f; As[r * 64 + f] = 0.0f; }
            }
        }
        __syncthreads();
        // ---- forward (nrc-train.comp:715-741) ----
        float* xin = XE0;
        float* xout = XE1;
#pragma unroll 1
        for (int l = 0; l < 5; l++) {
            copy16k(W, w.paramsT + l * 4096);
            __syncthreads();
            hidden_layer(xin, W, w.params + NRC_B_OFF(l), xout, As + (l + 1) * 4096);
            __syncthreads();
            float* t = xin; xin = xout; xout = t;
        }
        // xin == XE1 holds A5 feature-major; XE0 is free.
        if (tid < 192) {
            int r = tid & 63, j = tid >> 6;
            const float* w5 = w.params + NRC_W_OFF(5) + j * 64;
            float sum = 0.0f;
#pragma unroll 8
            for (int k = 0; k < 64; k++) sum = fmaf(xin[k * TILE + r], __ldg(w5 + k), sum);
            predS[r * 4 + j] = fmaxf(sum + __ldg(w.params + NRC_B_OFF(5) + j), 0.0f);
        }
        __syncthreads();
        // ---- loss and output error (nrc-train.comp:1002-1012) with the layer-5 gate of Backprop5 (:746-749) ----
        if (tid < 64) {
            int r = tid;
            float l = 0.0f;
            float e[3] = {0.0f, 0.0f, 0.0f};
            bool live = !dropS[r];
            if (live) {
                const float* a5 = As + 5 * 4096 + r * 64;
#pragma unroll
                for (int j = 0; j < 3; j++) {
                    float tg = fminf(__ldg(target + 3 * (base + r) + j), 1024.0f);
                    float err = predS[r * 4 + j] - tg;
                    l += err * err;
                    e[j] = (a5[j] > 0.0f) ? 2.0f * err : 0.0f;
                }
                l = (l / 3.0f) * scale;
            }
            e6S[r * 4 + 0] = e[0]; e6S[r * 4 + 1] = e[1]; e6S[r * 4 + 2] = e[2]; e6S[r * 4 + 3] = 0.0f;
            if (predOut && base + r < n) {
#pragma unroll
                for (int j = 0; j < 3; j++) predOut[3 * (base + r) + j] = live ? predS[r * 4 + j] : 0.0f;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) l += __shfl_xor_sync(0xffffffffu, l, o);
            if ((tid & 31) == 0) lossS[tid >> 5] = l;
        }
        __syncthreads();
        if (tid == 0) lossAcc += lossS[0] + lossS[1];
        // ---- layer 5: dW5, db5, and E5 = (W5^T e6) gated by A4 (Backprop5 :751-774 + Backprop4 gate :780-783) ----
        if (tid < 192) {
            int c = tid & 63, j = tid >> 6;
            const float* a5 = As + 5 * 4096;
            float s = 0.0f;
#pragma unroll 8
            for (int r = 0; r < 64; r++) s = fmaf(a5[r * 64 + c], e6S[r * 4 + j], s);
            partial[NRC_W_OFF(5) + j * 64 + c] += -s * scale;
        } else if (tid < 195) {
            int j = tid - 192;
            float s = 0.0f;
            for (int r = 0; r < 64; r++) s += e6S[r * 4 + j];
            partial[NRC_B_OFF(5) + j] += -s * scale;
        }
        {
            const float* w5 = w.params + NRC_W_OFF(5);
            const float* a4 = As + 4 * 4096;
            float ev[4][4];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                int r = 4 * tm + i;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    int c = 4 * tn + j;
                    float err = 0.0f;
                    err = fmaf(__ldg(w5 + c), e6S[r * 4 + 0], err);
                    err = fmaf(__ldg(w5 + 64 + c), e6S[r * 4 + 1], err);
                    err = fmaf(__ldg(w5 + 128 + c), e6S[r * 4 + 2], err);
                    ev[i][j] = (a4[r * 64 + c] > 0.0f) ? err : 0.0f;
                }
                *reinterpret_cast<float4*>(ES0 + r * 64 + 4 * tn) = make_float4(ev[i][0], ev[i][1], ev[i][2], ev[i][3]);
            }
#pragma unroll
            for (int j = 0; j < 4; j++)
                *reinterpret_cast<float4*>(XE0 + (4 * tn + j) * TILE + 4 * tm) = make_float4(ev[0][j], ev[1][j], ev[2][j], ev[3][j]);
        }
        __syncthreads();
        // ---- hidden layers 4..0 (Backprop4..Backprop0, nrc-train.comp:777-945) ----
        float* et = XE0; float* es = ES0;
        float* etn = XE1; float* esn = ES1;
#pragma unroll 1
        for (int l = 4; l >= 0; l--) {
            copy16k(W, w.params + NRC_W_OFF(l));   // row-major [out][in]: k-major for the back-data GEMM
            // weight gradient: dW_l[o][i] = -scale * sum_r E[r][o] * A_l[r][i]
            float acc[4][4];
            zero_acc(acc);
            gemm64(es, As + l * 4096, tm, tn, acc);
#pragma unroll
            for (int i = 0; i < 4; i++) {
                float4* dst = reinterpret_cast<float4*>(partial + NRC_W_OFF(l) + (4 * tm + i) * 64 + 4 * tn);
                float4 v = *dst;
                v.x += -acc[i][0] * scale; v.y += -acc[i][1] * scale; v.z += -acc[i][2] * scale; v.w += -acc[i][3] * scale;
                *dst = v;
            }
            if (tid < 64) {
                float s = 0.0f;
#pragma unroll 8
                for (int r = 0; r < 64; r++) s += es[r * 64 + tid];
                partial[NRC_B_OFF(l) + tid] += -s * scale;
            }
            __syncthreads();   // W staged
            // back-data: err[r][c] = sum_o E[r][o] * W_l[o][c]
            zero_acc(acc);
            gemm64(et, W, tm, tn, acc);
            const float* gate = (l > 0) ? (As + (l - 1) * 4096) : nullptr;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                int r = 4 * tm + i;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    int c = 4 * tn + j;
                    if (gate) acc[i][j] = (gate[r * 64 + c] > 0.0f) ? acc[i][j] : 0.0f;
                }
                *reinterpret_cast<float4*>(esn + r * 64 + 4 * tn) = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
            }
            if (gate) {
#pragma unroll
                for (int j = 0; j < 4; j++)
                    *reinterpret_cast<float4*>(etn + (4 * tn + j) * TILE + 4 * tm) = make_float4(acc[0][j], acc[1][j], acc[2][j], acc[3][j]);
            }
            __syncthreads();
            float* t1 = et; et = etn; etn = t1;
            float* t2 = es; es = esn; esn = t2;
        }
        // ---- dL/dx0 of the 32 hash features -> this sample's scratch row (mrhe_scatter.cuh, phase 1); zeros for dropped samples ----
        {
            int r = tid & 63, h = tid >> 6;
            size_t q = base + r;
            if (q < n) {
                const bool dead = dropS[r] != 0;
                float4* go = reinterpret_cast<float4*>(gAll + q * 32 + 8 * h);
                const float4* e = reinterpret_cast<const float4*>(es + r * 64 + 8 * h);
                go[0] = dead ? make_float4(0.0f, 0.0f, 0.0f, 0.0f) : e[0];
                go[1] = dead ? make_float4(0.0f, 0.0f, 0.0f, 0.0f) : e[1];
            }
        }
        __syncthreads();
    }
    // ---- hash-table gradient (BackpropMrhe, nrc-train.comp:947-983): phase 2, over all tiles of this CTA.
    // The histogram (parity preset, T <= NRC_HIST_MAX_T) takes over the activation buffers, which are dead now. ----
    mrhe_scatter_tiles<TILE>(pos, gAll, n, m, scale, histAll ? smem : nullptr, histAll ? histAll + (size_t)blockIdx.x * m.T : nullptr);
    if (tid == 0) partial[20995] = lossAcc;
}

// grad[i] += sum over CTAs (fixed order) of partial[b][i]; i == 20995 is the mseLoss accumulator.  Threads past the MLP arena add the
// CTAs' hash-gradient histograms (mrhe_scatter.cuh) into the head of the delta table: entry h receives the same sum in both features (A9).
__global__ void __launch_bounds__(256) k_reduce_partials(const float* __restrict__ partial, int nb, float* __restrict__ grad,
                                                         const float* __restrict__ hist, uint32_t T, float2* __restrict__ delta)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < NRC_MLP_PARAM_COUNT + 1) {
        float s = 0.0f;
        for (int b = 0; b < nb; b++) s += partial[(size_t)b * NRC_PARTIAL_STRIDE + i];
        grad[i] += s;
        return;
    }
    const uint32_t h = (uint32_t)i - NRC_HIST_FIRST;
    if (hist == nullptr || i < NRC_HIST_FIRST || h >= T) return;
    float s = 0.0f;
    for (int b = 0; b < nb; b++) s += hist[(size_t)b * T + h];
    float2 d = delta[h];
    d.x += s; d.y += s;
    delta[h] = d;
}

// Wt[l][in][out] = W[l][out][in] for the five hidden layers.
__global__ void __launch_bounds__(256) k_transpose_weights(const float* __restrict__ params, float* __restrict__ paramsT)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 5 * 4096) return;
    int l = i >> 12, rem = i & 4095, k = rem >> 6, c = rem & 63;
    paramsT[i] = params[l * 4096 + c * 64 + k];
}

// ---- host launchers -----------------------------------------------------------------------------------------

nrc_status launch_mlp_prepare(nrc_mlp* mlp)
{
    if (!mlp->dirtyT) return NRC_OK;
    NRC_LAUNCH(mlp->ctx, mlp->ctx->stream, K_TRANSPOSE_WEIGHTS,
               k_transpose_weights<<<(5 * 4096 + 255) / 256, 256, 0, mlp->ctx->stream>>>(mlp->params, mlp->paramsT));
    mlp->dirtyT = false;
    return NRC_OK;
}

nrc_status launch_infer_fp32(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const uint32_t* countDev, size_t n, float* rgb,
                             cudaStream_t st)
{
    if (n == 0) return NRC_OK;
    NRC_REQUIRE(m->F == 2 && m->L <= 16, "MLP kernels need featureCount == 2 and levelCount <= 16");
    NRC_TRY(launch_mlp_prepare(mlp));
    static bool attr[64] = {};   // per device: function attributes belong to the device's context
    size_t smem = 3 * 4096 * sizeof(float);
    if (first_on_device(mlp->ctx, attr)) {
        NRC_CUDA(cudaFuncSetAttribute(k_infer_fp32, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    size_t tiles = (n + TILE - 1) / TILE;
    size_t cap = (size_t)mlp->ctx->smCount * 4 * 4;   // 4 CTAs/SM resident, a few waves, grid-stride beyond
    unsigned blocks = (unsigned)(tiles < cap ? tiles : cap);
    MlpDev w{mlp->params, mlp->paramsT};
    NRC_LAUNCH(mlp->ctx, st, K_INFER_FP32, k_infer_fp32<<<blocks, NTHREADS, smem, st>>>(pos, dir, countDev, n, w, make_mrhe_dev(m), rgb));
    return NRC_OK;
}

nrc_status launch_reduce_partials(nrc_mlp* mlp, nrc_mrhe* m, int blocks, bool hist, cudaStream_t st)
{
    const uint32_t T = hist ? m->T : 0u;
    const unsigned threads = hist ? (NRC_HIST_FIRST + T) : (NRC_MLP_PARAM_COUNT + 1);
    NRC_LAUNCH(mlp->ctx, st, K_REDUCE_PARTIALS,
               k_reduce_partials<<<(threads + 255) / 256, 256, 0, st>>>(mlp->partial, blocks, mlp->grad, hist ? mlp->histPartial : nullptr, T,
                                                                        reinterpret_cast<float2*>(m->delta)));
    return NRC_OK;
}

// Scratch of one train launch: the per-CTA partial arenas, one dL/dx0 row per sample, and (parity preset with a table that fits shared
// memory) one histogram row per CTA.  Grown on demand; a growth waits for the device (the buffers may be in use by a queued launch).
nrc_status train_scratch(nrc_mlp* mlp, nrc_mrhe* m, size_t n, bool* useHist)
{
    const int sms = mlp->ctx->smCount;
    if (mlp->partialBlocks < sms) {
        NRC_CUDA(cudaDeviceSynchronize());
        if (mlp->partial) NRC_CUDA(cudaFree(mlp->partial));
        mlp->partial = nullptr;
        NRC_CUDA(cudaMalloc(&mlp->partial, (size_t)sms * NRC_PARTIAL_STRIDE * sizeof(float)));
        mlp->partialBlocks = sms;
    }
    if (mlp->gScratchRows < n) {
        NRC_CUDA(cudaDeviceSynchronize());
        if (mlp->gScratch) NRC_CUDA(cudaFree(mlp->gScratch));
        mlp->gScratch = nullptr; mlp->gScratchRows = 0;
        const size_t rows = n + n / 4 + 1024;
        NRC_CUDA(cudaMalloc(&mlp->gScratch, rows * 32 * sizeof(float)));
        mlp->gScratchRows = rows;
    }
    *useHist = !m->fixed && m->T <= NRC_HIST_MAX_T && (m->T % 4u) == 0u;
    if (*useHist && mlp->histPartialFloats < (size_t)sms * m->T) {
        NRC_CUDA(cudaDeviceSynchronize());
        if (mlp->histPartial) NRC_CUDA(cudaFree(mlp->histPartial));
        mlp->histPartial = nullptr; mlp->histPartialFloats = 0;
        NRC_CUDA(cudaMalloc(&mlp->histPartial, (size_t)sms * m->T * sizeof(float)));
        mlp->histPartialFloats = (size_t)sms * m->T;
    }
    return NRC_OK;
}

nrc_status launch_train_fp32(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const float* target, size_t n, float lossScale,
                             float* pred, cudaStream_t st, int* deferred)
{
    if (n == 0) return NRC_OK;
    NRC_REQUIRE(m->F == 2 && m->L <= 16, "MLP kernels need featureCount == 2 and levelCount <= 16");
    NRC_TRY(launch_mlp_prepare(mlp));
    static bool attr[64] = {};   // per device: function attributes belong to the device's context
    size_t smem = TRAIN_SMEM_FLOATS * sizeof(float);
    if (first_on_device(mlp->ctx, attr)) {
        NRC_CUDA(cudaFuncSetAttribute(k_train_fp32, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    bool useHist = false;
    NRC_TRY(train_scratch(mlp, m, n, &useHist));
    size_t tiles = (n + TILE - 1) / TILE;
    int blocks = (int)(tiles < (size_t)mlp->ctx->smCount ? tiles : (size_t)mlp->ctx->smCount);
    MlpDev w{mlp->params, mlp->paramsT};
    NRC_LAUNCH(mlp->ctx, st, K_TRAIN_FP32,
               k_train_fp32<<<blocks, NTHREADS, smem, st>>>(pos, dir, target, nullptr, n, lossScale, w, make_mrhe_dev(m), mlp->partial, pred,
                                                            mlp->gScratch, useHist ? mlp->histPartial : nullptr));
    if (deferred) { deferred[0] = blocks; deferred[1] = useHist ? 1 : 0; return NRC_OK; }
    return launch_reduce_partials(mlp, m, blocks, useHist, st);
}

// ---- allocation-free training entry point (no counterpart in the reference) ----------
// nrc_train_accumulate_static: Backprop() over up to NRC_STATIC_MAX_SAMPLES device-resident records with the fp32 kernel above, for
// callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it).  The scratch the train
// kernel needs -- per-CTA partial gradient arenas, one dL/dx0 row per sample, per-CTA hash-gradient histograms -- lives in ONE
// file-scope __device__ array of fixed size (about 3.4 MB, held by every GPU that loads the library until the process exits).
// Every call shares that array and it keeps the last call's contents: the kernel itself rewrites every word it later reads (each CTA
// zeroes its partial arena first, writes the dL/dx0 rows of all samples < n before phase 2 reads them, and stores its whole histogram
// row), so nothing has to be cleared between calls -- but two calls whose kernels overlap on the GPU (different streams / contexts of
// one device) race on it.  Calls on one context are stream-ordered and safe.
#define NRC_STATIC_MAX_BLOCKS 16
#define NRC_STATIC_MAX_SAMPLES 8192
#define NRC_STATIC_MAX_T 16384
#define NRC_STATIC_PARTIAL_FLOATS (NRC_STATIC_MAX_BLOCKS * NRC_PARTIAL_STRIDE)
#define NRC_STATIC_G_FLOATS (NRC_STATIC_MAX_SAMPLES * 32)
#define NRC_STATIC_HIST_FLOATS (NRC_STATIC_MAX_BLOCKS * NRC_STATIC_MAX_T)
#define NRC_STATIC_FLOATS (NRC_STATIC_PARTIAL_FLOATS + NRC_STATIC_G_FLOATS + NRC_STATIC_HIST_FLOATS)
static __device__ __align__(256) float g_staticTrainScratch[NRC_STATIC_FLOATS];

// In host code the array's name is only a stand-in: the device address comes from cudaGetSymbolAddress, once per device, with that
// device current.
#define NRC_STATIC_MAX_DEVICES 64
static float* g_staticTrainScratchDev[NRC_STATIC_MAX_DEVICES] = {nullptr};

extern "C" size_t nrc_train_accumulate_static_max_samples(void) { return NRC_STATIC_MAX_SAMPLES; }

extern "C" nrc_status nrc_train_accumulate_static(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const float* target, size_t n,
                                                  float lossScale, float* pred)
{
    NRC_REQUIRE(mlp && m && pos && dir && target, "null argument");
    NRC_REQUIRE(mlp->ctx == m->ctx, "mlp and mrhe belong to different contexts");
    NRC_REQUIRE(n <= NRC_STATIC_MAX_SAMPLES, "nrc_train_accumulate_static: more than 8192 samples (the static scratch is sized for that); "
                                             "use nrc_train_accumulate or split the batch");
    NRC_REQUIRE(m->F == 2 && m->L <= 16, "MLP kernels need featureCount == 2 and levelCount <= 16");
    NRC_REQUIRE(nrc_is_device_ptr(pos) && nrc_is_device_ptr(dir) && nrc_is_device_ptr(target) && (!pred || nrc_is_device_ptr(pred)),
                "nrc_train_accumulate_static takes device memory only (staging host buffers would allocate)");
    if (n == 0) return NRC_OK;
    nrc_ctx* ctx = mlp->ctx;
    NRC_REQUIRE(ctx->device >= 0 && ctx->device < NRC_STATIC_MAX_DEVICES, "device index out of range for the static scratch");
    NRC_CUDA(cudaSetDevice(ctx->device));
    if (!g_staticTrainScratchDev[ctx->device]) {
        float* p = nullptr;
        NRC_CUDA(cudaGetSymbolAddress((void**)&p, g_staticTrainScratch));
        g_staticTrainScratchDev[ctx->device] = p;
    }
    float* scratch = g_staticTrainScratchDev[ctx->device];
    float* partial = scratch;
    float* gRows = scratch + NRC_STATIC_PARTIAL_FLOATS;
    float* hist = gRows + NRC_STATIC_G_FLOATS;
    const bool useHist = !m->fixed && m->T <= NRC_STATIC_MAX_T && (m->T % 4u) == 0u;   // larger tables / fixed preset: global atomics
    cudaStream_t st = ctx->stream;
    NRC_TRY(launch_mlp_prepare(mlp));
    static bool attr[64] = {};   // per device: function attributes belong to the device's context
    const size_t smem = TRAIN_SMEM_FLOATS * sizeof(float);
    if (first_on_device(mlp->ctx, attr)) {
        NRC_CUDA(cudaFuncSetAttribute(k_train_fp32, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    const size_t tiles = (n + TILE - 1) / TILE;
    const int blocks = (int)(tiles < NRC_STATIC_MAX_BLOCKS ? tiles : NRC_STATIC_MAX_BLOCKS);
    MlpDev w{mlp->params, mlp->paramsT};
    NRC_LAUNCH(ctx, st, K_TRAIN_FP32,
               k_train_fp32<<<blocks, NTHREADS, smem, st>>>(pos, dir, target, nullptr, n, lossScale, w, make_mrhe_dev(m), partial, pred, gRows,
                                                            useHist ? hist : nullptr));
    const uint32_t T = useHist ? m->T : 0u;
    const unsigned threads = useHist ? (NRC_HIST_FIRST + T) : (NRC_MLP_PARAM_COUNT + 1);
    NRC_LAUNCH(ctx, st, K_REDUCE_PARTIALS,
               k_reduce_partials<<<(threads + 255) / 256, 256, 0, st>>>(partial, blocks, mlp->grad, useHist ? hist : nullptr, T,
                                                                        reinterpret_cast<float2*>(m->delta)));
    return NRC_OK;
}
