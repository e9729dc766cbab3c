// mlp_tc.cu -- fused MRHE/OneBlob encode + 64-wide MLP on the 5th-generation tensor cores (tcgen05, sm_100a).
//
// SURVEY.md 8a rows a9-a11 (EncodeRay + Forward, nrc-forward.frag:503-534) as one kernel per query batch:
//
//   CTA tile = 128 queries = the 128 TMEM lanes.  256 threads: thread (row, half) encodes half of the 64 network
//   inputs of its row (8 MRHE levels + 16 OneBlob bins) straight into the A operand in shared memory (bf16).
//   Each hidden layer is D[128x64] = A[128x64] * W_l^T: four tcgen05.mma.kind::f16 (M=128, N=64, K=16) issued by ONE
//   thread, fp32 accumulators in TMEM, completion signalled through tcgen05.commit -> mbarrier.  The epilogue reads
//   the accumulator with tcgen05.ld (warp w owns TMEM lanes 32*(w%4).., column half w/4), adds the fp32 bias, applies
//   ReLU, rounds to bf16 and writes the next layer's A operand back to the same shared-memory tile.  All six weight
//   matrices stay resident in shared memory (43 KB bf16) for the lifetime of the CTA; ~61 KB of shared memory and 64
//   TMEM columns per CTA let three CTAs share an SM, so one CTA's gathers/epilogue overlap another's MMAs.
//
// Operand layout (both operands K-major, no swizzle = "interleave" canonical layout of the UMMA shared-memory
// descriptor): a core matrix is 8 rows x 16 bytes (8 bf16 along K), stored as 128 contiguous bytes;
//     byte(row, k) = (k / 8) * (ROWS * 16) + row * 16 + (k % 8) * 2
// so LBO (core matrices adjacent in K) = ROWS*16 and SBO (adjacent 8-row groups) = 128.  A thread writing the 16-byte
// chunk of its own row gives 512 contiguous bytes per warp: conflict-free, no swizzle arithmetic.
//
// Precision: operands are bf16 (8-bit mantissa), accumulation fp32, biases fp32.  Against the fp32 oracle the stated
// bound is 3e-2 of the largest output magnitude (tests/test_gpu_tc.py); the fp32 CUDA-core kernels (mlp_fp32.cu) stay
// the 1e-3 parity path.
#include <cuda_bf16.h>
#include <stdlib.h>
#include <string.h>
#include "nrc_internal.cuh"
#include "strict.cuh"

#include "tc_ptx.cuh"


// all four operand images from the packed fp32 arena (the padding of the layer-5 images stays zero from the allocation's memset)
__global__ void __launch_bounds__(256) k_pack_bf16(const float* __restrict__ params, unsigned short* __restrict__ img)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < NRC_MLP_WEIGHT_COUNT) tc_store_weight(img, i, params[i]);
}

// ---- shared-memory map ----------------------------------------------------------------------------------------------------
#define TC_SMEM_W 0                               // bf16 weight image, 43 008 B
#define TC_SMEM_A (TC_W_ELEMS * 2)                // A tile: 8 K-chunks x 128 rows x 16 B = 16 384 B
#define TC_SMEM_BIAS (TC_SMEM_A + 16384)          // 323 fp32 biases (336 slots)
#define TC_SMEM_BAR (TC_SMEM_BIAS + 336 * 4)      // mbarrier (8 B) + TMEM base slot (4 B)
#define TC_SMEM_BYTES (TC_SMEM_BAR + 16)

// 8 consecutive network inputs of one row as one 16-byte chunk: bf16 (SPLIT = false) or the three bf16 parts hi / mid / lo of the fp32
// values into three tiles of the same layout (the split-precision forward pass of the train kernel)
template <bool SPLIT>
__device__ __forceinline__ void tc_store_chunk(uint4* hi, uint4* mid, uint4* lo, int index, const float* o)
{
    if (!SPLIT) {
        hi[index] = make_uint4(pack_bf16(o[0], o[1]), pack_bf16(o[2], o[3]), pack_bf16(o[4], o[5]), pack_bf16(o[6], o[7]));
        return;
    }
    uint32_t h[4], m[4], l[4];
#pragma unroll
    for (int j = 0; j < 4; j++) split2_bf16x3(o[2 * j], o[2 * j + 1], h[j], m[j], l[j]);
    hi[index] = make_uint4(h[0], h[1], h[2], h[3]);
    mid[index] = make_uint4(m[0], m[1], m[2], m[3]);
    lo[index] = make_uint4(l[0], l[1], l[2], l[3]);
}

// The network inputs of one tile (EncodeRay: fwd.frag:483-501, train.comp:703-713) into the A-operand image, by PARTS threads per query row
// (2: the 256-thread kernels; 4: the 512-thread train kernel).  Neighbouring lanes (2j, 2j+1) share a row: lane parity = the x SIDE of
// the cell.  Each lane gathers the four corners of its side (both features) and multiplies them by its side's factor; one
// shuffle per corner pair hands the partner the product it needs, and each lane finishes ONE feature (parity = feature) through the y
// and z stages.  The products, the sums and their order are those of mrhe_blend() (the first stage is a sum of two products and a + b
// = b + a), so the encoded values are bit-identical to the one-thread form -- with half the hashes and blends per thread, and, because
// the two sides of a corner pair are neighbouring table entries (mrhe_level_indices_side), half the 128-byte lines per load instruction:
// the encoders are bound by L1 lines per load (~2 cycles per line, 32 lines for 32 scattered lanes), not by instructions.
// A second round of shuffles pairs the two features of a level for the bf16x2 words: lane 0 of a pair assembles the chunks of the first
// half of the pair's levels, lane 1 those of the second half.  OneBlob: PARTS slices of the 32 bins.
//   SPLIT: three bf16 parts hi / mid / lo into three tiles (the train kernel's split-precision forward pass)
//   GATES (training): gate0[8 r + c] bit j = (network input 8 c + j > 0) on the fp32 value -- the layer-0 ReLU gate of the reference
//   (the error of layer 1 is gated by the raw input features).  It cannot be read off the bf16 operand: OneBlob bins a few
//   sigma away from the angle are fp32 denormals (> 0) that round to +0 in bf16, for ~20 % of all directions.  dropS[r] = row r is past
//   n or has a NaN among its first five inputs (train.comp:993-999).
//   ZERO (training, PARTS = 2): a dropped row's inputs are stored as zeros right here (the drop decision travels to the pair's other
//   lane by one more shuffle); without it the caller zeroes the rows after a barrier.
// Every thread of the CTA (of the calling warps) must call this (full-warp shuffles).
// Against one thread per half row (round 1's form): 104 M instead of 90.9 M warp instructions per frame in k_infer_tc (both lanes of a pair
// set up the level) and still the faster one in the frame (2.760 / 2.762 against 2.775 / 2.779 ms in alternating runs), 0.128 -> 0.109 ms
// for 2^18 scattered queries, k_encode_tiles 0.169 -> 0.139 ms at 2^19 (profiles/r02y_ncu_summary.md, r02t_train_tc.md).
template <int PARTS, bool SPLIT, bool GATES, bool ZERO = false>
__device__ __forceinline__ void tc_encode_pair(unsigned char* aTile, unsigned char* aMid, unsigned char* aLo, int tid, size_t tileBase, size_t n,
                                               const float* __restrict__ pos, const float* __restrict__ dir, const MrheDev& m,
                                               unsigned char* gate0, int* dropS)
{
    constexpr int LV = 32 / PARTS;   // levels per lane pair: 16 or 8
    constexpr int H = LV / 2;        // levels whose chunks one lane assembles
    constexpr int BC = 4 / PARTS;    // OneBlob chunks (8 bins each) per thread
    static_assert(!ZERO || (PARTS == 2 && GATES), "ZERO: both lanes of a row's pair are in one warp");
    const int side = tid & 1, r = (tid >> 1) & (TC_ROWS - 1), lh = tid >> 8;
    const size_t q = tileBase + (size_t)r;
    const bool live = q < n;
    f3 p = mk3(0.0f, 0.0f, 0.0f), d = mk3(0.0f, 0.0f, 1.0f);
    if (live) {
        p = mk3(__ldg(pos + 3 * q), __ldg(pos + 3 * q + 1), __ldg(pos + 3 * q + 2));
        d = mk3(__ldg(dir + 3 * q), __ldg(dir + 3 * q + 1), __ldg(dir + 3 * q + 2));
    }
    uint4* chunk = reinterpret_cast<uint4*>(aTile);   // chunk(kc, row) = chunk[kc * 128 + row]
    uint4* chunkM = reinterpret_cast<uint4*>(aMid);
    uint4* chunkL = reinterpret_cast<uint4*>(aLo);
    const f3 sky = mk3(m.skySize[0], m.skySize[1], m.skySize[2]);
    const f3 np = mrhe_norm_pos(p, sky);
    float o[LV];   // feature `side` of levels LV * lh + i
#pragma unroll
    for (int i = 0; i < LV; i++) {
        const uint32_t level = LV * lh + i;
        float2 f[4];
        f3 lerp = mk3(0.0f, 0.0f, 0.0f);
#pragma unroll
        for (int k = 0; k < 4; k++) f[k] = make_float2(0.0f, 0.0f);
        if (live && level < m.L) {
            uint32_t idx[4];
            mrhe_level_indices_side(p, np, m.res[level], m.T, side, idx, lerp, m.fixed != 0);
            lerp = mrhe_blend_factors(lerp, m.fixed != 0);
            const float2* tab = reinterpret_cast<const float2*>(m.tables) + (size_t)level * m.T;
#pragma unroll
            for (int k = 0; k < 4; k++) f[k] = __ldg(tab + idx[k]);
        }
        // mrhe_blend(): z_k = f[x0 corner k] * (1 - l.z) + f[x1 corner k] * l.z, then y, then x
        const float w = side ? lerp.z : NRC_SUB(1.0f, lerp.z);
        float z[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const float a0 = NRC_MUL(f[k].x, w), a1 = NRC_MUL(f[k].y, w);
            const float got = __shfl_xor_sync(0xffffffffu, side ? a0 : a1, 1);   // the partner's product for MY feature
            z[k] = side ? NRC_ADD(got, a1) : NRC_ADD(a0, got);
        }
        const float iy = NRC_SUB(1.0f, lerp.y), ix = NRC_SUB(1.0f, lerp.x);
        const float y0 = NRC_ADD(NRC_MUL(z[0], iy), NRC_MUL(z[2], lerp.y));
        const float y1 = NRC_ADD(NRC_MUL(z[1], iy), NRC_MUL(z[3], lerp.y));
        o[i] = NRC_ADD(NRC_MUL(y0, ix), NRC_MUL(y1, lerp.x));
    }
    // pair the features: lane 0 keeps levels [0, H), lane 1 levels [H, LV)
    float f0[H], f1[H];
#pragma unroll
    for (int j = 0; j < H; j++) {
        const float got = __shfl_xor_sync(0xffffffffu, side ? o[j] : o[H + j], 1);
        f0[j] = side ? got : o[j];
        f1[j] = side ? o[H + j] : got;
    }
    bool drop = false;   // ZERO: this row is past n or has a NaN among its first five inputs (lane 0 of the pair holds them)
    if (ZERO) {
        int d0 = (!live || (f0[0] != f0[0]) | (f1[0] != f1[0]) | (f0[1] != f0[1]) | (f1[1] != f1[1]) | (f0[2] != f0[2])) ? 1 : 0;
        drop = __shfl_sync(0xffffffffu, d0, (tid & 31) & ~1) != 0;
        if (side == 0) dropS[r] = drop ? 1 : 0;
    }
#pragma unroll
    for (int g = 0; g < H / 4; g++) {
        const int ci = lh * (LV / 4) + side * (H / 4) + g;   // chunk = levels 4 ci .. 4 ci + 3
        float v[8];
#pragma unroll
        for (int li = 0; li < 4; li++) { v[2 * li] = f0[4 * g + li]; v[2 * li + 1] = f1[4 * g + li]; }
        if (GATES) {
            if (!ZERO && ci == 0) dropS[r] = (!live || (v[0] != v[0]) | (v[1] != v[1]) | (v[2] != v[2]) | (v[3] != v[3]) | (v[4] != v[4])) ? 1 : 0;
            unsigned gm = 0u;
#pragma unroll
            for (int j = 0; j < 8; j++) gm |= (v[j] > 0.0f ? 1u : 0u) << j;
            gate0[8 * r + ci] = (unsigned char)gm;
        }
        if (ZERO && drop) {
#pragma unroll
            for (int j = 0; j < 8; j++) v[j] = 0.0f;
        }
        tc_store_chunk<SPLIT>(chunk, chunkM, chunkL, ci * TC_ROWS + r, v);
    }
    // OneBlob (train.comp:344-365): slice bq of the 32 bins
    const int bq = side + 2 * lh;
    float ang;
    if (BC * bq < 2) {
        ang = NRC_ADD(NRC_DIV(nrc_atan2(d.z, d.x), NRC_PI), 0.5f);
    } else {
        float lenxz = NRC_SQRT(NRC_ADD(NRC_MUL(d.x, d.x), NRC_MUL(d.z, d.z)));
        ang = NRC_ADD(NRC_DIV(nrc_atan2(lenxz, d.y), NRC_PI), 0.5f);
    }
    const float sigma = 0.25f;
    const float term1 = NRC_DIV(1.0f, NRC_MUL(sigma, NRC_SQRT(NRC_MUL(2.0f, NRC_PI))));
#pragma unroll
    for (int g = 0; g < BC; g++) {
        const int bc = BC * bq + g;
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            float fi = (float)(8 * (bc & 1) + i);
            float t2 = NRC_DIV(NRC_SUB(fi, ang), sigma);
            v[i] = live ? NRC_MUL(term1, nrc_exp(NRC_MUL(NRC_MUL(-0.5f, t2), t2))) : 0.0f;
        }
        if (GATES) {
            unsigned gb = 0u;
#pragma unroll
            for (int i = 0; i < 8; i++) gb |= (v[i] > 0.0f ? 1u : 0u) << i;
            gate0[8 * r + 4 + bc] = (unsigned char)gb;
        }
        if (ZERO && drop) {
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = 0.0f;
        }
        tc_store_chunk<SPLIT>(chunk, chunkM, chunkL, (4 + bc) * TC_ROWS + r, v);
    }
}

__global__ void __launch_bounds__(TC_THREADS, 3) k_infer_tc(const float* __restrict__ pos, const float* __restrict__ dir,
                                                            const uint32_t* __restrict__ countDev, size_t nMax,
                                                            const __nv_bfloat16* __restrict__ wimg, const float* __restrict__ params, MrheDev m,
                                                            float* __restrict__ rgb)
{
    extern __shared__ __align__(128) unsigned char smem[];
    unsigned char* aTile = smem + TC_SMEM_A;
    float* biasS = reinterpret_cast<float*>(smem + TC_SMEM_BIAS);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + TC_SMEM_BAR);
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(smem + TC_SMEM_BAR + 8);

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int quad = warp & 3, half = warp >> 2;
    const int row = quad * 32 + lane;

    // ---- one-time setup: weights + biases -> smem, mbarrier, TMEM ----
    {
        const uint4* src = reinterpret_cast<const uint4*>(wimg);
        uint4* dst = reinterpret_cast<uint4*>(smem + TC_SMEM_W);
        for (int i = tid; i < TC_W_ELEMS * 2 / 16; i += TC_THREADS) dst[i] = __ldg(src + i);
        for (int i = tid; i < 336; i += TC_THREADS) biasS[i] = (i < 323) ? __ldg(params + NRC_B_OFF(0) + i) : 0.0f;
    }
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(tmemSlot, TC_TMEM_COLS);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmemBase = *tmemSlot;
    const uint32_t aAddr = smem_u32(aTile), wAddr = smem_u32(smem + TC_SMEM_W);
    uint32_t phase = 0;

    size_t n = nMax;
    if (countDev) { size_t c = *countDev; n = c < nMax ? c : nMax; }
    const size_t tiles = (n + TC_ROWS - 1) / TC_ROWS;
    for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const size_t q = tile * TC_ROWS + row;   // the row this thread reads out of the accumulator
        const bool live = q < n;
        tc_encode_pair<2, false, false>(aTile, nullptr, nullptr, tid, tile * TC_ROWS, n, pos, dir, m, nullptr, nullptr);
        fence_async_smem();
        tc_fence_before();
        __syncthreads();
#pragma unroll 1
        for (int l = 0; l < 6; l++) {
            if (tid == 0) {
                tc_fence_after();
                const uint32_t nOut = (l < 5) ? 64u : 16u;
                const uint32_t idesc = umma_idesc(TC_ROWS, nOut);
                const uint32_t wBase = wAddr + (uint32_t)l * 8192u;          // layer 5 starts at 5*8192 = 40960
                const uint32_t wLbo = nOut * 16u;                             // 64 rows -> 1024 B, 16 rows -> 256 B
#pragma unroll
                for (int ks = 0; ks < 4; ks++) {
                    uint64_t ad = umma_desc(aAddr + ks * 2 * (TC_ROWS * 16), TC_ROWS * 16, 128);
                    uint64_t bd = umma_desc(wBase + ks * 2 * wLbo, wLbo, 128);
                    umma_bf16(tmemBase, ad, bd, idesc, ks > 0 ? 1u : 0u);
                }
                umma_commit(bar);
            }
            mbar_wait(bar, phase);
            phase ^= 1u;
            tc_fence_after();
            if (l < 5) {
                uint32_t v[32];
                tmem_ld32(tmemBase + ((uint32_t)(quad * 32) << 16) + (uint32_t)(half * 32), v);
                tmem_ld_wait();
                // + bias as packed adds (FADD2), ReLU and the rounding to bf16 in one cvt.rn.relu.bf16x2.f32 per pair: one instruction per
                // value instead of 2.5 (the epilogues were half of this kernel's instructions); the same values as fmaxf + cvt.rn
                const float4* b = reinterpret_cast<const float4*>(biasS + l * 64 + half * 32);
                uint4* chunk = reinterpret_cast<uint4*>(aTile);
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const float4 b0 = b[2 * c], b1 = b[2 * c + 1];
                    float x[8];
                    add2_bits(x[0], x[1], v[8 * c], v[8 * c + 1], b0.x, b0.y);
                    add2_bits(x[2], x[3], v[8 * c + 2], v[8 * c + 3], b0.z, b0.w);
                    add2_bits(x[4], x[5], v[8 * c + 4], v[8 * c + 5], b1.x, b1.y);
                    add2_bits(x[6], x[7], v[8 * c + 6], v[8 * c + 7], b1.z, b1.w);
                    chunk[(4 * half + c) * TC_ROWS + row] = make_uint4(pack_relu_bf16(x[0], x[1]), pack_relu_bf16(x[2], x[3]),
                                                                       pack_relu_bf16(x[4], x[5]), pack_relu_bf16(x[6], x[7]));
                }
                fence_async_smem();
            } else if (half == 0) {
                uint32_t v[4];
                tmem_ld4(tmemBase + ((uint32_t)(quad * 32) << 16), v);
                tmem_ld_wait();
                if (live) {
#pragma unroll
                    for (int j = 0; j < 3; j++) rgb[3 * q + j] = fmaxf(__uint_as_float(v[j]) + biasS[320 + j], 0.0f);
                }
            }
            tc_fence_before();
            __syncthreads();
        }
    }
    tc_fence_after();
    if (warp == 0) tmem_dealloc(tmemBase, TC_TMEM_COLS);
}

// ---- host launchers -----------------------------------------------------------------------------------------------------------
nrc_status tc_prepare(nrc_mlp* mlp, cudaStream_t st)
{
    if (!mlp->paramsBf16) {
        NRC_CUDA(cudaMalloc(&mlp->paramsBf16, TC_IMG_ELEMS * sizeof(uint16_t)));
        NRC_CUDA(cudaMemsetAsync(mlp->paramsBf16, 0, TC_IMG_ELEMS * sizeof(uint16_t), st));
        mlp->dirtyBf16 = true;
    }
    if (mlp->dirtyBf16) {
        NRC_LAUNCH(mlp->ctx, st, K_PACK_BF16,
                   k_pack_bf16<<<(NRC_MLP_WEIGHT_COUNT + 255) / 256, 256, 0, st>>>(mlp->params, mlp->paramsBf16));
        mlp->dirtyBf16 = false;
    }
    return NRC_OK;
}

nrc_status launch_infer_tc(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const uint32_t* countDev, size_t n, float* rgb,
                           cudaStream_t st)
{
    if (n == 0) return NRC_OK;
    NRC_REQUIRE(m->F == 2 && m->L <= 16, "MLP kernels need featureCount == 2 and levelCount <= 16");
    NRC_TRY(tc_prepare(mlp, st));
    static bool attr[64] = {};   // per device: function attributes belong to the device's context
    if (first_on_device(mlp->ctx, attr)) {
        NRC_CUDA(cudaFuncSetAttribute(k_infer_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM_BYTES));
    }
    size_t tiles = (n + TC_ROWS - 1) / TC_ROWS;
    size_t cap = (size_t)mlp->ctx->smCount * 3;   // persistent: 3 CTAs per SM, grid-stride over the tiles (1 / 2 / 4 per SM: 2.87 / 2.78 / 2.78 ms per frame against 2.77)
    unsigned blocks = (unsigned)(tiles < cap ? tiles : cap);
    NRC_LAUNCH(mlp->ctx, st, K_INFER_TC,
               k_infer_tc<<<blocks, TC_THREADS, TC_SMEM_BYTES, st>>>(pos, dir, countDev, n, reinterpret_cast<const __nv_bfloat16*>(mlp->paramsBf16),
                                                                     mlp->params, make_mrhe_dev(m), rgb));
    return NRC_OK;
}

// ---- split inference: encode -> bf16 operand-image tiles in HBM -> k_mlp_stream (mlp_stream.cu) ---------------------------
// One block per 128-query tile, no shared memory, no TMEM: a pure gather kernel at full occupancy.  The tile image in
// global memory is byte for byte the A-operand image k_infer_tc builds in shared memory (chunk(kc,row) = 16 B at
// (kc*128 + row)*16), so the MLP kernel brings a tile in with ONE 16 KB bulk copy and a warp's stores are 512 contiguous bytes.
__global__ void __launch_bounds__(TC_THREADS) k_encode_tiles(const float* __restrict__ pos, const float* __restrict__ dir,
                                                             const uint32_t* __restrict__ countDev, size_t nMax, MrheDev m,
                                                             unsigned char* __restrict__ featTiles)
{
    size_t n = nMax;
    if (countDev) { size_t c = *countDev; n = c < nMax ? c : nMax; }
    const size_t tile = blockIdx.x;
    if (tile * TC_ROWS >= n) return;
    tc_encode_pair<2, false, false>(featTiles + tile * 16384, nullptr, nullptr, (int)threadIdx.x, tile * TC_ROWS, n, pos, dir, m, nullptr, nullptr);
}

nrc_status launch_encode_tiles(nrc_mrhe* m, const float* pos, const float* dir, const uint32_t* countDev, size_t n, void* featTiles,
                               cudaStream_t st)
{
    if (n == 0) return NRC_OK;
    NRC_REQUIRE(m->F == 2 && m->L <= 16, "MLP kernels need featureCount == 2 and levelCount <= 16");
    size_t tiles = (n + TC_ROWS - 1) / TC_ROWS;
    NRC_LAUNCH(m->ctx, st, K_ENCODE_TILES,
               k_encode_tiles<<<(unsigned)tiles, TC_THREADS, 0, st>>>(pos, dir, countDev, n, make_mrhe_dev(m), (unsigned char*)featTiles));
    return NRC_OK;
}

nrc_status launch_infer_split(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const uint32_t* countDev, size_t n, float* rgb,
                              cudaStream_t st)
{
    if (n == 0) return NRC_OK;
    size_t bytes = ((n + TC_ROWS - 1) / TC_ROWS) * 16384;
    if (mlp->featBytes < bytes) {
        if (mlp->featTiles) { NRC_CUDA(cudaDeviceSynchronize()); NRC_CUDA(cudaFree(mlp->featTiles)); mlp->featTiles = nullptr; mlp->featBytes = 0; }
        NRC_CUDA(cudaMalloc(&mlp->featTiles, bytes));
        mlp->featBytes = bytes;
    }
    NRC_TRY(launch_encode_tiles(m, pos, dir, countDev, n, mlp->featTiles, st));
    return launch_mlp_stream(mlp, mlp->featTiles, countDev, n, rgb, st);
}

// ==================================================================================================================
// Training on the tensor cores: EncodeRay + Forward + Backprop (train.comp:703-741, 743-1028) for a tile of 128 samples.
//
//   forward   A_{l+1} = relu(A_l W_l^T + b_l)                  tcgen05.mma M=128 (samples) N=64 K=64 in SPLIT PRECISION: activations and
//                                                              weights are kept as three bf16 parts (x = hi + mid + lo) and a K step is
//                                                              six MMAs (hi*hi, hi*mid, mid*hi, hi*lo, lo*hi, mid*mid) into one fp32
//                                                              accumulator -- the fp32 product to ~2^-24.  Why the forward pass: the
//                                                              ReLU gates of the WHOLE backward pass are read off the forward
//                                                              activations, and a plain bf16 forward flips enough of them (|z| below
//                                                              the rounding noise) to put 8-30 % of error on the hash-table gradient,
//                                                              whose corner weights are unbounded so that a few samples
//                                                              dominate an entry.  Measured by emulation (scripts/tc_split_emul.py):
//                                                              bf16 forward 12-30 %, two-part split 0.4-6 %, three-part split <= 0.75 %
//                                                              = the floor set by the bf16 backward pass; making the last back-data GEMM
//                                                              precise instead changes nothing.  The tensor pipe idles in this kernel
//                                                              (gathers and epilogues bound it), so 6x the forward MMAs are free.
//                                                              The hi part of every A_l stays in shared memory for the backward pass
//                                                              (gates, dW operands); mid / lo live only until the next layer has read them.
//   back-data E_l     = gate(E_{l+1} W_l)                      M=128 N=64 K=64, bf16, B = the TRANSPOSED weight image
//   dW_l, db_l        = E_{l+1}^T [A_l | 1]                    M=128 (64 out rows; the other 64 lanes are never read) N=64 (+16 for db: a
//                                                              constant ones tile) K=128 (the samples).  E_{l+1} and A_l are read WHERE
//                                                              THEY LIE as MN-major operands (the feature index is the contiguous one
//                                                              in the K-major tiles the other GEMMs use): no transposed copies.  The
//                                                              accumulators of all five hidden layers live in TMEM (5 x 80 columns)
//                                                              across ALL tiles of the CTA and are read once at the end -> per-CTA
//                                                              partials, summed in fixed order by k_reduce_partials.  No atomics on the
//                                                              MLP parameters.
//   layer 5 (64 -> 3): forward and back-data on the tensor cores (N = 16 / K = 16 padded), dW5/db5 on CUDA cores.
//   gates, loss, clamp, NaN guard: the reference's.
//   hash-table gradient: dL/dx0 of every sample goes to a global scratch row; after the last tile the CTA scatters its samples into a
//   shared-memory histogram (mrhe_scatter.cuh) -- no global atomics in the parity preset.
// TMEM: columns 0-63 forward/back-data accumulator, 64 + 80*l .. dW_l | db_l (464 of 512 columns).
// ==================================================================================================================
#include "mrhe_scatter.cuh"

#define TT_WF 0                                    // forward weight stages hi | mid | lo, 8 KB each
#define TT_WB 24576                                // transposed weight stage (8 KB)
// WS = the warp-specialised form (below): a second network-input set that the encoder warps fill while the first is in use
template <bool WS>
struct TT {
    static constexpr uint32_t ACT1 = 32768;                        // A_1..A_5 (hi parts), 16 KB each, K-major (K = feature)
    static constexpr uint32_t SET0 = ACT1 + 5 * 16384;             // the network inputs A_0 of a tile: hi | lo | mid, 16 KB each.  hi stays for the
                                                                   //   backward pass (gates, dW_0 operand); mid / lo then carry the parts of every
                                                                   //   layer's input, and lo becomes the E tile (K-major A operand of the back-data
                                                                   //   GEMMs; read MN-major by the dW GEMMs, whose M = 128 runs 64 feature rows past
                                                                   //   it into mid: lanes nobody reads, inside the allocation)
    static constexpr uint32_t SET1 = SET0 + 49152;                 // WS only
    static constexpr uint32_t ONES = WS ? SET1 + 49152 : SET1;     // 128 samples x 16 features, MN-major: feature 0 = 1 (bf16), the rest 0 (4 KB): db = E^T 1
    static constexpr uint32_t BIAS = ONES + 4096;                  // 336 floats
    static constexpr uint32_t E6 = BIAS + 336 * 4;                 // [128][4] fp32 output errors
    static constexpr uint32_t PRED = E6 + 2048;                    // [128][4] fp32 predictions
    static constexpr uint32_t DROP = PRED + 2048;                  // [2][128] int (one per input set)
    static constexpr uint32_t LOSS = DROP + 1024;                  // [8] float
    static constexpr uint32_t GATE0 = LOSS + 64;                   // [2][128][8] bytes: (input > 0) bits of the 64 fp32 network inputs (layer-0 gate)
    static constexpr uint32_t BAR = GATE0 + 2048;
    static constexpr uint32_t BYTES = BAR + 32;
    static constexpr uint32_t HIST = ACT1;                         // after the last tile: the hash-gradient histogram (up to 128 KiB = A_1..A_5 + set 0)
    static_assert(BYTES <= 227 * 1024, "k_train_tc shared memory");
    static_assert(HIST + NRC_HIST_MAX_T * 4 <= SET1, "histogram overlaps live buffers");
};
#define TT_DW_COL(l) (64u + 80u * (uint32_t)(l))

// named barriers of the warp-specialised form (0 is __syncthreads)
#define TT_BAR_COMPUTE 1
#define TT_BAR_FULL 2      // + set: the encoder warps have filled input set `set`
#define TT_BAR_EMPTY 4     // + set: the compute warps are done with input set `set`
__device__ __forceinline__ void named_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }
__device__ __forceinline__ void named_arrive(int id, int count) { asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory"); }

// Weight staging by ONE thread with bulk copies (cp.async.bulk, completion counted on an mbarrier): the other 255 threads go straight
// to their epilogue instead of stalling on L2 loads (13 % of the kernel's stall samples in profiles/r02t were the staging loop).
// the three parts of layer l's forward weights into the three stages
__device__ __forceinline__ void tt_bulk_fwd(unsigned char* smem, const __nv_bfloat16* __restrict__ wimg, int l, uint64_t* barW)
{
    const uint32_t bytes = (l < 5) ? 8192u : 2048u;
    const int off = (l < 5) ? l * 4096 : 5 * 4096;
    mbar_expect_tx(barW, 3u * bytes);
    bulk_load(smem + TT_WF, wimg + off, bytes, barW);
    bulk_load(smem + TT_WF + 8192, wimg + TC_IMG_MID + off, bytes, barW);
    bulk_load(smem + TT_WF + 16384, wimg + TC_IMG_LO + off, bytes, barW);
}
// layer l's transposed weights into the backward stage
__device__ __forceinline__ void tt_bulk_bwd(unsigned char* smem, const __nv_bfloat16* __restrict__ wimg, int l, uint64_t* barW)
{
    const uint32_t bytes = (l < 5) ? 8192u : 2048u;
    mbar_expect_tx(barW, bytes);
    bulk_load(smem + TT_WB, wimg + TC_IMG_T + ((l < 5) ? l * 4096 : 5 * 4096), bytes, barW);
}

// PARTS threads per sample row (2: 256 threads, each 32 accumulator columns / 8 levels; 4: 512 threads, 16 columns / 4 levels): the
// scatter's shared-memory atomics are latency-bound and halve with twice the warps.
// WS (with PARTS = 2): warp-specialised, 512 threads.  Threads 0-255 run the tile (forward, backward, dW5) exactly as in the plain form;
// threads 256-511 ENCODE THE NEXT TILE meanwhile into the other of two input sets (a tile's encode is a third of its time and is bound
// by gathered sectors from L2, the rest by issue slots and shared memory: they overlap).  Hand-over by named barriers: FULL[set]
// (encoders arrive, compute warps wait), EMPTY[set] (compute warps arrive after their last read of the set, encoders wait before they
// overwrite it two tiles later).  512-thread CTAs want most of a register file and wait for a nearly empty SM: both 512-thread forms are
// for batches of many tiles per CTA; inside a frame (one tile per CTA, beside the tracing kernels) the plain 256-thread form runs.
template <int PARTS, bool WS>
__global__ void __launch_bounds__(WS ? 512 : 128 * PARTS, 1) k_train_tc(const float* __restrict__ pos, const float* __restrict__ dir,
                                                            const float* __restrict__ target, size_t n, float scale,
                                                            const __nv_bfloat16* __restrict__ wimg, const float* __restrict__ params, MrheDev m,
                                                            float* __restrict__ partialAll, float* __restrict__ predOut, float* gAll,
                                                            float* __restrict__ histAll)
{
    extern __shared__ __align__(128) unsigned char smem[];
    static_assert(!WS || PARTS == 2, "the warp-specialised form computes with two threads per row");
    using M = TT<WS>;
    float* biasS = reinterpret_cast<float*>(smem + M::BIAS);
    float* e6S = reinterpret_cast<float*>(smem + M::E6);
    float* predS = reinterpret_cast<float*>(smem + M::PRED);
    float* lossS = reinterpret_cast<float*>(smem + M::LOSS);
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem + M::BAR);
    uint32_t* tmemSlot = reinterpret_cast<uint32_t*>(smem + M::BAR + 8);
    uint64_t* barW = reinterpret_cast<uint64_t*>(smem + M::BAR + 16);   // weight stages filled (bulk copies of thread 0)
    constexpr int CT = 128 * PARTS;              // threads that run the tile
    constexpr int THREADS = WS ? 2 * CT : CT;    // all threads of the CTA
    constexpr int COLS = 64 / PARTS;       // accumulator columns per thread
    constexpr int CH = COLS / 8;           // 16-byte chunks (8 features) per thread

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int quad = warp & 3, part = warp >> 2;   // TMEM lanes 32 quad .. +31 are this warp's; part = which COLS columns of them
    const int row = quad * 32 + lane;
    float* partial = partialAll + (size_t)blockIdx.x * NRC_PARTIAL_STRIDE;

    // ---- one-time setup ----
    for (int i = tid; i < NRC_PARTIAL_STRIDE; i += THREADS) partial[i] = 0.0f;
    for (int i = tid; i < 4096 / 16; i += THREADS) reinterpret_cast<uint4*>(smem + M::ONES)[i] = make_uint4(0u, 0u, 0u, 0u);
    for (int i = tid; i < 336; i += THREADS) biasS[i] = (i < 323) ? __ldg(params + NRC_B_OFF(0) + i) : 0.0f;
    if (tid == 0) {
        mbar_init(bar, 1);
        mbar_init(barW, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) tmem_alloc(tmemSlot, 512);
    __syncthreads();
    if (tid < 128) reinterpret_cast<uint4*>(smem + M::ONES)[tid] = make_uint4(0x3F80u, 0u, 0u, 0u);   // [1 0 .. 0] per sample: column 64 = db
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmemBase = *tmemSlot;
    const uint32_t sBase = smem_u32(smem);
    uint32_t phase = 0, phaseW = 0;   // phaseW: thread 0 only
    float lossAcc = 0.0f;        // thread 0
    float dw5Acc = 0.0f;         // threads < 192: dW5[j][c]; threads 192..194: db5[j]

    const size_t tiles = (n + TC_ROWS - 1) / TC_ROWS;
    bool firstTile = true;
    uint32_t kTile = 0;   // this CTA's k-th tile: input set k & 1 in the warp-specialised form
    if (WS && tid >= CT) {
        // ---- encoder warps: tile k into set k & 1, one tile ahead of the compute warps ----
        for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x, kTile++) {
            const uint32_t set = kTile & 1u, so = M::SET0 + set * 49152u;
            if (kTile >= 2u) named_sync(TT_BAR_EMPTY + (int)set, THREADS);   // the tile that used this set two tiles ago is done with it
            tc_encode_pair<2, true, true, true>(smem + so, smem + so + 32768, smem + so + 16384, tid - CT, tile * TC_ROWS, n, pos, dir, m,
                                                smem + M::GATE0 + set * 1024u, reinterpret_cast<int*>(smem + M::DROP + set * 512u));
            fence_async_smem();
            named_arrive(TT_BAR_FULL + (int)set, THREADS);
        }
    } else
    for (size_t tile = blockIdx.x; tile < tiles; tile += gridDim.x, firstTile = false, kTile++) {
        const size_t q = tile * TC_ROWS + row;   // the sample whose accumulator row this thread reads
        const bool inRange = q < n;
        const uint32_t set = WS ? (kTile & 1u) : 0u;
        const uint32_t hiOff = M::SET0 + set * 49152u, loOff = hiOff + 16384u, midOff = hiOff + 32768u;   // lo doubles as the E tile
        const int* dropS = reinterpret_cast<const int*>(smem + M::DROP + set * 512u);
        const unsigned char* gate0S = smem + M::GATE0 + set * 1024u;
        auto actOff = [&](int l) -> uint32_t { return l == 0 ? hiOff : M::ACT1 + (uint32_t)(l - 1) * 16384u; };   // hi part of A_l
        if (tid == 0) tt_bulk_fwd(smem, wimg, 0, barW);   // the forward stages are idle since the last tile's layer 5
        bool dropped;
        if (WS) {
            named_sync(TT_BAR_FULL + (int)set, THREADS);   // the encoder warps have filled this set (dropped rows already zero)
            dropped = dropS[row] != 0;
        } else {
            // ---- encode into A_0 = hi + mid + lo (EncodeRay, train.comp:703-713) + NaN guard (train.comp:993-999) ----
            tc_encode_pair<PARTS, true, true, false>(smem + hiOff, smem + midOff, smem + loOff, tid, tile * TC_ROWS, n, pos, dir, m,
                                                     smem + M::GATE0, reinterpret_cast<int*>(smem + M::DROP));
            __syncthreads();
            dropped = dropS[row] != 0;
            if (dropped) {   // a dropped sample contributes nothing: zero its inputs (its error rows are zeroed below)
                for (int c = CH * part; c < CH * part + CH; c++) {
                    reinterpret_cast<uint4*>(smem + hiOff)[c * TC_ROWS + row] = make_uint4(0u, 0u, 0u, 0u);
                    reinterpret_cast<uint4*>(smem + midOff)[c * TC_ROWS + row] = make_uint4(0u, 0u, 0u, 0u);
                    reinterpret_cast<uint4*>(smem + loOff)[c * TC_ROWS + row] = make_uint4(0u, 0u, 0u, 0u);
                }
            }
            fence_async_smem();
            tc_fence_before();
            __syncthreads();
        }
        // ---- forward (train.comp:715-741), split precision ----
#pragma unroll 1
        for (int l = 0; l < 6; l++) {
            if (tid == 0) {
                mbar_wait(barW, phaseW);   // this layer's weights have landed
                phaseW ^= 1u;
                tc_fence_after();
                const uint32_t nOut = (l < 5) ? 64u : 16u;
                const uint32_t idesc = umma_idesc(TC_ROWS, nOut);
                const uint32_t wLbo = nOut * 16u;
                const uint32_t aH = sBase + actOff(l), aM = sBase + midOff, aL = sBase + loOff;
                const uint32_t wH = sBase + TT_WF, wM = wH + 8192u, wL = wH + 16384u;
#pragma unroll
                for (int ks = 0; ks < 4; ks++) {
                    const uint32_t ao = (uint32_t)ks * 4096u, wo = (uint32_t)ks * 2u * wLbo;
                    // smallest partial products first
                    umma_bf16(tmemBase, umma_desc(aM + ao, 2048, 128), umma_desc(wM + wo, wLbo, 128), idesc, ks > 0 ? 1u : 0u);
                    umma_bf16(tmemBase, umma_desc(aL + ao, 2048, 128), umma_desc(wH + wo, wLbo, 128), idesc, 1u);
                    umma_bf16(tmemBase, umma_desc(aH + ao, 2048, 128), umma_desc(wL + wo, wLbo, 128), idesc, 1u);
                    umma_bf16(tmemBase, umma_desc(aM + ao, 2048, 128), umma_desc(wH + wo, wLbo, 128), idesc, 1u);
                    umma_bf16(tmemBase, umma_desc(aH + ao, 2048, 128), umma_desc(wM + wo, wLbo, 128), idesc, 1u);
                    umma_bf16(tmemBase, umma_desc(aH + ao, 2048, 128), umma_desc(wH + wo, wLbo, 128), idesc, 1u);
                }
                umma_commit(bar);
            }
            mbar_wait(bar, phase);
            phase ^= 1u;
            tc_fence_after();
            // the forward stages are free again: fetch the next layer's weights (or the first backward stage) behind the epilogue
            if (tid == 0) {
                if (l < 5) tt_bulk_fwd(smem, wimg, l + 1, barW);
                else tt_bulk_bwd(smem, wimg, 5, barW);     // W5^T: 64 rows x K 16
            }
            if (l < 5) {
                uint32_t v[COLS];
                tmem_ld_cols<COLS>(tmemBase + ((uint32_t)(quad * 32) << 16) + (uint32_t)(part * COLS), v);
                tmem_ld_wait();
                const float* b = biasS + l * 64 + part * COLS;
                uint4* cH = reinterpret_cast<uint4*>(smem + M::ACT1 + l * 16384);
                uint4* cM = reinterpret_cast<uint4*>(smem + midOff);
                uint4* cL = reinterpret_cast<uint4*>(smem + loOff);
#pragma unroll
                for (int c = 0; c < CH; c++) {
                    float x[8];
#pragma unroll
                    for (int j = 0; j < 8; j++) x[j] = fmaxf(__uint_as_float(v[8 * c + j]) + b[8 * c + j], 0.0f);
                    tc_store_chunk<true>(cH, cM, cL, (CH * part + c) * TC_ROWS + row, x);
                }
            } else if (part == 0) {
                // ---- output, loss, output error with the layer-5 gate (train.comp:1002-1012, 746-749) ----
                uint32_t v[4];
                tmem_ld4(tmemBase + ((uint32_t)(quad * 32) << 16), v);
                tmem_ld_wait();
                const unsigned short* a5 = reinterpret_cast<const unsigned short*>(smem + M::ACT1 + 4 * 16384 + row * 16);   // features 0..7 of this row
                float lsum = 0.0f, e[3];
#pragma unroll
                for (int j = 0; j < 3; j++) {
                    float y = fmaxf(__uint_as_float(v[j]) + biasS[320 + j], 0.0f);
                    float err = 0.0f;
                    e[j] = 0.0f;
                    if (!dropped) {
                        float tg = fminf(__ldg(target + 3 * q + j), 1024.0f);
                        err = y - tg;
                        lsum += err * err;
                        bool gate = (a5[j] & 0x7fffu) != 0u && (a5[j] & 0x8000u) == 0u;     // hi part > 0  <=>  value > 0
                        e[j] = gate ? 2.0f * err : 0.0f;
                    }
                    predS[row * 4 + j] = dropped ? 0.0f : y;
                    e6S[row * 4 + j] = e[j];
                }
                e6S[row * 4 + 3] = 0.0f;
                if (predOut && inRange) {
#pragma unroll
                    for (int j = 0; j < 3; j++) predOut[3 * q + j] = predS[row * 4 + j];
                }
                // E_6 as the A operand of the layer-5 back-data GEMM (K = 16: two chunks); the lo part that lived here is dead
                uint4* echunk = reinterpret_cast<uint4*>(smem + loOff);
                echunk[row] = make_uint4(pack_bf16(e[0], e[1]), pack_bf16(e[2], 0.0f), 0u, 0u);
                echunk[TC_ROWS + row] = make_uint4(0u, 0u, 0u, 0u);
                lsum = (lsum / 3.0f) * scale;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) lsum += __shfl_xor_sync(0xffffffffu, lsum, o);
                if (lane == 0) lossS[quad] = lsum;
            }
            fence_async_smem();
            tc_fence_before();
            if (WS) named_sync(TT_BAR_COMPUTE, CT); else __syncthreads();
        }
        if (tid == 0) lossAcc += (lossS[0] + lossS[1]) + (lossS[2] + lossS[3]);
        // ---- dW5, db5 on the CUDA cores (Backprop5, train.comp:751-774): 192 + 3 threads, accumulated over the CTA's tiles ----
        if (tid < 192) {
            int c = tid & 63, j = tid >> 6;
            const unsigned char* a5 = smem + M::ACT1 + 4 * 16384 + (c >> 3) * 2048 + (c & 7) * 2;
            float s = 0.0f;
#pragma unroll 8
            for (int r = 0; r < TC_ROWS; r++)
                s = fmaf(__bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(a5 + r * 16)), e6S[r * 4 + j], s);
            dw5Acc += s;
        } else if (tid < 195) {
            int j = tid - 192;
            float s = 0.0f;
            for (int r = 0; r < TC_ROWS; r++) s += e6S[r * 4 + j];
            dw5Acc += s;
        }
        // ---- backward (Backprop5..Backprop0, train.comp:743-945).  Step L computes dX_L = E_{L+1} W_L into the main
        // accumulator (and, for L <= 4, dW_L | db_L += E_{L+1}^T [A_L | 1] into its own TMEM columns); its epilogue gates dX_L
        // with A_{L-1} into E_L -- K-major for the next back-data GEMM and transposed for the next dW GEMM --
        // and stages what step L-1 needs: W_{L-1}^T and [A_{L-1} | 1]^T. ----
#pragma unroll 1
        for (int L = 5; L >= 0; L--) {
            if (tid == 0) {
                mbar_wait(barW, phaseW);   // W_L^T has landed
                phaseW ^= 1u;
                tc_fence_after();
                if (L <= 4) {
                    // K = sample: E_{L+1} and A_L are read where they lie, as MN-major operands (8 features of one sample per 16 bytes,
                    // 8 samples per 128-byte core matrix: leading offset 128 B between groups of 8 samples, stride 2048 B between
                    // groups of 8 features).  M = 128 reads 64 features past the E tile: those accumulator lanes are never read.
                    const uint32_t idw = umma_idesc(128, 64) | UMMA_A_MN | UMMA_B_MN, idb = umma_idesc(128, 16) | UMMA_A_MN | UMMA_B_MN;
#pragma unroll
                    for (int ks = 0; ks < 8; ks++) {
                        const uint64_t eT = umma_desc(sBase + loOff + ks * 256, 128, 2048);
                        umma_bf16(tmemBase + TT_DW_COL(L), eT, umma_desc(sBase + actOff(L) + ks * 256, 128, 2048), idw,
                                  (firstTile && ks == 0) ? 0u : 1u);
                        umma_bf16(tmemBase + TT_DW_COL(L) + 64u, eT, umma_desc(sBase + M::ONES + ks * 256, 128, 2048), idb,
                                  (firstTile && ks == 0) ? 0u : 1u);
                    }
                }
                const uint32_t idx = umma_idesc(TC_ROWS, 64);
                const int ksteps = (L == 5) ? 1 : 4;
                for (int ks = 0; ks < ksteps; ks++)
                    umma_bf16(tmemBase, umma_desc(sBase + loOff + ks * 4096, 2048, 128), umma_desc(sBase + TT_WB + ks * 2048, 1024, 128), idx,
                              ks > 0 ? 1u : 0u);
                umma_commit(bar);
            }
            mbar_wait(bar, phase);
            phase ^= 1u;
            tc_fence_after();
            uint32_t v[COLS];
            tmem_ld_cols<COLS>(tmemBase + ((uint32_t)(quad * 32) << 16) + (uint32_t)(part * COLS), v);
            tmem_ld_wait();
            if (L >= 1) {
                if (tid == 0) tt_bulk_bwd(smem, wimg, L - 1, barW);
                const uint4* aprev = reinterpret_cast<const uint4*>(smem + actOff(L - 1));
                uint4* echunk = reinterpret_cast<uint4*>(smem + loOff);
                // gates: A_{L-1} > 0.  Hidden activations: the sign of the hi part; the network inputs (L == 1): the exact fp32 signs
#pragma unroll
                for (int c = 0; c < CH; c++) {
                    const uint32_t g0 = gate0S[8 * row + CH * part + c];
                    uint4 a = aprev[(CH * part + c) * TC_ROWS + row];       // A_{L-1}[row][COLS part + 8c .. +7]
                    uint32_t aw[4] = {a.x, a.y, a.z, a.w};
                    uint32_t ew[4];
#pragma unroll
                    for (int j2 = 0; j2 < 4; j2++) {
                        unsigned short alo = (unsigned short)(aw[j2] & 0xffffu), ahi = (unsigned short)(aw[j2] >> 16);
                        float elo = __uint_as_float(v[8 * c + 2 * j2]), ehi = __uint_as_float(v[8 * c + 2 * j2 + 1]);
                        bool openLo = (alo & 0x7fffu) != 0u && (alo & 0x8000u) == 0u, openHi = (ahi & 0x7fffu) != 0u && (ahi & 0x8000u) == 0u;
                        if (L == 1) { openLo = (g0 >> (2 * j2)) & 1u; openHi = (g0 >> (2 * j2 + 1)) & 1u; }
                        if (!openLo) elo = 0.0f;
                        if (!openHi) ehi = 0.0f;
                        ew[j2] = pack_bf16(elo, ehi);
                    }
                    echunk[(CH * part + c) * TC_ROWS + row] = make_uint4(ew[0], ew[1], ew[2], ew[3]);
                }
            } else if (part * COLS < 32 && inRange) {
                // dX_0: the hash encoding's gradient lives in features 0..31 -> this sample's scratch row (mrhe_scatter.cuh, phase 1)
                float4* go = reinterpret_cast<float4*>(gAll + q * 32 + part * COLS);
#pragma unroll
                for (int j = 0; j < (COLS < 32 ? COLS : 32) / 4; j++)
                    go[j] = dropped ? make_float4(0.0f, 0.0f, 0.0f, 0.0f)
                                    : make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                                                  __uint_as_float(v[4 * j + 3]));
            }
            fence_async_smem();
            tc_fence_before();
            if (WS) named_sync(TT_BAR_COMPUTE, CT); else __syncthreads();
        }
        if (WS) named_arrive(TT_BAR_EMPTY + (int)set, THREADS);   // nothing of this tile reads its input set any more
    }
    if (WS) __syncthreads();   // encoder and compute warps meet again: the histogram below overwrites the activations
    // ---- hash-table gradient (BackpropMrhe, train.comp:947-983): phase 2, over all tiles of this CTA ----
    mrhe_scatter_tiles<TC_ROWS, THREADS>(pos, gAll, n, m, scale, histAll ? reinterpret_cast<float*>(smem + M::HIST) : nullptr,
                                histAll ? histAll + (size_t)blockIdx.x * m.T : nullptr);
    // ---- gradients out: TMEM -> this CTA's partial arena (negative gradient times ONE_OVER_PIXEL_COUNT, A8) ----
    tc_fence_after();
    if (tid < CT && quad < 2) {
        const int o = quad * 32 + lane;   // output feature = TMEM lane
#pragma unroll 1
        for (int l = 0; l < 5; l++) {
            uint32_t v[COLS];
            tmem_ld_cols<COLS>(tmemBase + ((uint32_t)(quad * 32) << 16) + TT_DW_COL(l) + (uint32_t)(part * COLS), v);
            tmem_ld_wait();
            float4* dst = reinterpret_cast<float4*>(partial + NRC_W_OFF(l) + o * 64 + part * COLS);
#pragma unroll
            for (int j = 0; j < COLS / 4; j++)
                dst[j] = make_float4(-__uint_as_float(v[4 * j]) * scale, -__uint_as_float(v[4 * j + 1]) * scale,
                                     -__uint_as_float(v[4 * j + 2]) * scale, -__uint_as_float(v[4 * j + 3]) * scale);
            if (part == PARTS - 1) {
                uint32_t b[4];
                tmem_ld4(tmemBase + ((uint32_t)(quad * 32) << 16) + TT_DW_COL(l) + 64u, b);
                tmem_ld_wait();
                partial[NRC_B_OFF(l) + o] = -__uint_as_float(b[0]) * scale;
            }
        }
    }
    if (tid < 192) partial[NRC_W_OFF(5) + (tid >> 6) * 64 + (tid & 63)] = -dw5Acc * scale;
    else if (tid < 195) partial[NRC_B_OFF(5) + (tid - 192)] = -dw5Acc * scale;
    if (tid == 0) partial[20995] = lossAcc;
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmemBase, 512);
}

nrc_status launch_reduce_partials(nrc_mlp* mlp, nrc_mrhe* m, int blocks, bool hist, cudaStream_t st);   // mlp_fp32.cu
nrc_status train_scratch(nrc_mlp* mlp, nrc_mrhe* m, size_t n, bool* useHist);                            // mlp_fp32.cu

nrc_status launch_train_tc(nrc_mlp* mlp, nrc_mrhe* m, const float* pos, const float* dir, const float* target, size_t n, float lossScale,
                           float* pred, cudaStream_t st, int* deferred)
{
    // the opt-in fixed preset (per-feature errors, per-level gradient tables) trains on the fp32 CUDA-core kernel
    if (m->fixed) return launch_train_fp32(mlp, m, pos, dir, target, n, lossScale, pred, st, deferred);
    if (n == 0) return NRC_OK;
    NRC_REQUIRE(m->F == 2 && m->L <= 16, "MLP kernels need featureCount == 2 and levelCount <= 16");
    NRC_TRY(tc_prepare(mlp, st));
    static bool attr[64] = {};   // per device: function attributes belong to the device's context
    if (first_on_device(mlp->ctx, attr)) {
        NRC_CUDA(cudaFuncSetAttribute(k_train_tc<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TT<false>::BYTES));
        NRC_CUDA(cudaFuncSetAttribute(k_train_tc<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TT<false>::BYTES));
        NRC_CUDA(cudaFuncSetAttribute(k_train_tc<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TT<true>::BYTES));
    }
    bool useHist = false;
    NRC_TRY(train_scratch(mlp, m, n, &useHist));
    size_t tiles = (n + TC_ROWS - 1) / TC_ROWS;
    int blocks = (int)(tiles < (size_t)mlp->ctx->smCount ? tiles : (size_t)mlp->ctx->smCount);
    // The form: a frame's batch (10^4 samples, one tile per CTA) runs beside the tracing kernels, where a 512-thread CTA would wait for a
    // nearly empty SM: plain, 256 threads.  A stand-alone batch with several tiles per CTA: warp-specialised (encoders one tile ahead).
    // NRC_TRAIN_TC_FORM=256 / 512 / ws forces one (read at every launch: tests switch it).
    int form = tiles >= 4 * (size_t)mlp->ctx->smCount ? 2 : 0;
    if (const char* e = getenv("NRC_TRAIN_TC_FORM")) form = !strcmp(e, "ws") ? 2 : (atoi(e) == 512 ? 1 : 0);
    const __nv_bfloat16* wimg = reinterpret_cast<const __nv_bfloat16*>(mlp->paramsBf16);
    float* hist = useHist ? mlp->histPartial : nullptr;
    const MrheDev md = make_mrhe_dev(m);
    if (form == 2)
        NRC_LAUNCH(mlp->ctx, st, K_TRAIN_TC, (k_train_tc<2, true><<<blocks, 512, TT<true>::BYTES, st>>>(pos, dir, target, n, lossScale, wimg, mlp->params, md,
                                                                                                       mlp->partial, pred, mlp->gScratch, hist)));
    else if (form == 1)
        NRC_LAUNCH(mlp->ctx, st, K_TRAIN_TC, (k_train_tc<4, false><<<blocks, 512, TT<false>::BYTES, st>>>(pos, dir, target, n, lossScale, wimg, mlp->params, md,
                                                                                                         mlp->partial, pred, mlp->gScratch, hist)));
    else
        NRC_LAUNCH(mlp->ctx, st, K_TRAIN_TC, (k_train_tc<2, false><<<blocks, 256, TT<false>::BYTES, st>>>(pos, dir, target, n, lossScale, wimg, mlp->params, md,
                                                                                                         mlp->partial, pred, mlp->gScratch, hist)));
    if (deferred) { deferred[0] = blocks; deferred[1] = useHist ? 1 : 0; return NRC_OK; }
    return launch_reduce_partials(mlp, m, blocks, useHist, st);
}
